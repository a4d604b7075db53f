#!/bin/bash
# A/B at N GPUs: cacheable vs write-combined pinned host staging (MGAR_WC_HOST) - end-to-end and copy-only GB/s
N=$1
for wc in 0 1; do
  MGAR_WC_HOST=$wc python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$wc bench.py --gpus $N --steps 20 --warmup 5 --quick --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
e=d['e2e']; print('N=$N wc=$wc value', round(d['value']), 'e2e', round(e['value']), 'h2d', e['h2d_GBps_per_gpu'], 'h2d_only', e['h2d_only_GBps_per_gpu'], e['h2d_only_GBps_all_gpus'])"
done
