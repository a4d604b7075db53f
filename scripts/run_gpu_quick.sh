#!/bin/bash
# quick GPU check: MLP / pipeline parity tests + a short bench
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_pipeline_gpu.py tests/test_ref_gpu_parity.py -q -m gpu -p no:cacheprovider -x ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_quick.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print(d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()}, 'e2e', d['e2e']['value'], 'lat', d['latency_1frame_ms'])
PY
tail -3 gpurun_out/bench.err
