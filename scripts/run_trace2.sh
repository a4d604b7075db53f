#!/bin/bash
# timelines of the tcgen05 MLP kernel from trace builds: scripts/run_trace2.sh libmgar_tr_a.so libmgar_tr_b.so ...
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
for lib in "$@"; do
  MGAR_LIB=$PWD/multimodal_gar_b200/$lib timeout 300 python bench.py --steps 1 --warmup 1 --quick --no-cpu-baseline > /dev/null 2> gpurun_out/trace.err
  grep tc16trace gpurun_out/trace.err | awk '$3!=-1' | tail -32 > gpurun_out/trace_$lib.txt; wc -l gpurun_out/trace_$lib.txt
done
