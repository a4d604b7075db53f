#!/bin/bash
# one ncu --set full capture of each of the five kernels of a step (after the plain command has exited 0)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sa_mlp_tc|ball_query|fps_|pool_from_mask|box_mask" -s 5 -c 5 \
    -o gpurun_out/prof_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
