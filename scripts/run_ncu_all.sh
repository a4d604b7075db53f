#!/bin/bash
# one ncu --set full capture of each kernel of a step (after the plain command has exited 0), then the launch list
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --quick --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:"sa_mlp_tc16|fps_bq|pool_from_mask|box_mask|build_live_units|order_clouds|frame_grid" -s 16 -c 8 \
    -o gpurun_out/prof_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
