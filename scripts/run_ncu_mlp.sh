#!/bin/bash
# one ncu --set full capture of the SA-MLP kernel (after the plain command has exited 0)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_K:-sa_mlp_tc}" -s ${NCU_S:-1} -c ${NCU_C:-1} \
    -o gpurun_out/${NCU_O:-prof_mlp} -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log
