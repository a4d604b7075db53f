#!/bin/bash
# scripts/run_exp_env.sh "ENV=val ..." lib.so ["ENV=val ..." lib.so ...] : quick bench per (environment, library) pair
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
while [ $# -ge 2 ]; do
  envs=$1; lib=$2; shift 2
  env $envs MGAR_LIB=$PWD/multimodal_gar_b200/$lib timeout 200 python bench.py --quick --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/exp.err | python -c "
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$envs $lib', round(d['value']), d['ms_per_step'], {k: v['ms'] for k, v in d['stages'].items()})
" || tail -3 gpurun_out/exp.err
done
