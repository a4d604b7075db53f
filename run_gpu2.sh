#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_parity_gpu.py -q -m gpu -p no:cacheprovider -k "tensor_core or sa_m" -x > gpurun_out/pytest_tc.log 2>&1; echo "tc rc=$?"
tail -40 gpurun_out/pytest_tc.log
