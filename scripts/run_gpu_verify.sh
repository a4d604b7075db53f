#!/bin/bash
# GPU verification round: parity tests, smoke, bench (both arms), training-step timing
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python scripts/time_train.py > gpurun_out/train_timing.log 2>&1; echo "train rc=$?"; tail -2 gpurun_out/train_timing.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
timeout 120 python scripts/time_dafm.py > gpurun_out/dafm_timing.log 2>&1; echo "dafm rc=$?"; tail -1 gpurun_out/dafm_timing.log
timeout 120 python scripts/diag_dafm.py > gpurun_out/dafm_diag.log 2>&1; tail -1 gpurun_out/dafm_diag.log
timeout 300 python scripts/time_config3.py > gpurun_out/config3_timing.log 2>&1; echo "config3 rc=$?"; tail -1 gpurun_out/config3_timing.log
