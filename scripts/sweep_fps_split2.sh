#!/bin/bash
# split threshold x resident CTAs per SM of the periodic-cloud kernel
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
for c in ${CTAS:-10}; do for v in ${SPLITS:-320 336 352 368 384}; do
  MGAR_FPS_BQ_CTAS=$c MGAR_FPS_BQ_SPLIT=$v timeout 200 python bench.py --quick --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/split.err | python -c "
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ctas=$c split=$v', round(d['value']), d['ms_per_step'], d['stages']['fps_ball_query']['ms'])
" || tail -3 gpurun_out/split.err
done; done
