#!/bin/bash
# A/B of one environment switch on the bench: scripts/run_ab.sh VAR A B
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
for v in "$2" "$3" "$2" "$3"; do
  env "$1=$v" timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_$v.json 2> gpurun_out/ab.err || tail -5 gpurun_out/ab.err
  python - "$1=$v" gpurun_out/ab_$v.json <<'PY'
import json, sys
d = json.load(open(sys.argv[2]))
print(sys.argv[1], round(d['value']), d['ms_per_step'], {k: v['ms'] for k, v in d['stages'].items()}, 'e2e', round(d['e2e']['value']))
PY
done
