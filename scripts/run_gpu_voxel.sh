#!/bin/bash
# voxel query: parity tests against the reference extension + golden generation for the CPU suite + timing
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_voxel_query_gpu.py tests/test_roiaware_gpu.py tests/test_pipeline_gpu.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_voxel.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/pytest_voxel.log
timeout 300 python tests/golden/make_ref_voxel_query_golden.py 2>&1 | tail -8
timeout 300 python scripts/time_voxel_query.py 2>&1 | tail -4
