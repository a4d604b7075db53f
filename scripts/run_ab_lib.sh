#!/bin/bash
# A/B of two builds of the library on the bench: scripts/run_ab_lib.sh prev.so cur.so  (alternating, 2 runs each)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
i=0
for lib in "$1" "$2" "$1" "$2"; do
  i=$((i+1))
  MGAR_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_$i.json 2> gpurun_out/ab.err || tail -5 gpurun_out/ab.err
  python - "$lib" gpurun_out/ab_$i.json <<'PY'
import json, sys
d = json.load(open(sys.argv[2]))
print(sys.argv[1], round(d['value']), d['ms_per_step'], {k: v['ms'] for k, v in d['stages'].items()}, 'e2e', round(d['e2e']['value']), d['gpu_launches'])
PY
done
