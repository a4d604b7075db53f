#!/bin/bash
# pointnet2_batch surface: parity tests against the reference extension
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_batch_ops_gpu.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_batch.log 2>&1; echo "pytest rc=$?"
tail -40 gpurun_out/pytest_batch.log
