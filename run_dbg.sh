#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for d in 0 2 16 18 22; do
MGAR_TC_DBG=$d timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu-baseline > gpurun_out/bq.json 2> gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bq.json')); print('dbg',$d,'sa_mlp ms', d['stages']['sa_mlp']['ms'])"
done
