#!/bin/bash
# first GPU round: reference goldens, parity tests, smoke, bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
python tests/golden/make_ref_gpu_golden.py gpurun_out/ref_gpu.npz > gpurun_out/golden.log 2>&1; echo "golden rc=$?"
cp gpurun_out/ref_gpu.npz tests/golden/ref_gpu.npz 2>/dev/null
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/pytest_gpu.log
timeout 600 python -m pytest tests/test_oracle_pin.py -q -p no:cacheprovider > gpurun_out/pytest_pin.log 2>&1; echo "pin rc=$?"; tail -5 gpurun_out/pytest_pin.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
