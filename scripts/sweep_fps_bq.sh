#!/bin/bash
# A/B of the fused FPS + ball query kernel's geometry and residency cap on the bench step (GPU box)
for cfg in "2,8 4" "2,8 6" "2,8 8" "2,8 10" "2,8 14" "4,4 2" "4,4 3" "4,4 4" "4,4 6" "8,2 2" "8,2 3"; do
  set -- $cfg
  MGAR_FPS_BQ_GEOM=$1 MGAR_FPS_BQ_CTAS=$2 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1 ctas=$2', round(d['value']), {k:v['ms'] for k,v in d['stages'].items()})"
done
