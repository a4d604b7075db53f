#!/bin/bash
# fused FPS + ball query: split threshold sweep on the bench step (MGAR_FPS_BQ_SPLIT=0: the single two-warp kernel)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
for v in "$@"; do
  MGAR_FPS_BQ_SPLIT=$v timeout 200 python bench.py --quick --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/split.err | python -c "
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('split=$v', round(d['value']), d['ms_per_step'], {k: v['ms'] for k, v in d['stages'].items()})
" || tail -3 gpurun_out/split.err
done
