#!/bin/bash
# bench (quick) on the regular library and on the timing-only experiment builds libmgar_exp<n>.so (scripts/build_exp.sh)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
for lib in libmgar_lidar.so "$@"; do
  MGAR_LIB=$PWD/multimodal_gar_b200/$lib timeout 300 python bench.py --quick --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/exp.err | python -c "
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib', round(d['value']), d['ms_per_step'], {k: v['ms'] for k, v in d['stages'].items()})
" || tail -3 gpurun_out/exp.err
done
