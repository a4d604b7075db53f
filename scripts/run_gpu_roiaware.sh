#!/bin/bash
# RoIAwarePool3d: parity tests against the reference extension + golden generation for the CPU suite
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_roiaware_gpu.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_roiaware.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/pytest_roiaware.log
timeout 300 python tests/golden/make_ref_roiaware_golden.py 2>&1 | tail -8
timeout 300 python scripts/time_roiaware.py 2>&1 | tail -8
