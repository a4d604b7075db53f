#!/bin/bash
for c in 3 4 6 8 12 16; do
  MGAR_FPS_BQ_CTAS=$c python bench.py --steps 20 --warmup 5 --quick --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('dup ctas=$c', round(d['value']), {k:v['ms'] for k,v in d['stages'].items()})"
done
