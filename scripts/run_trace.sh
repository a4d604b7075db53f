#!/bin/bash
# debug: timeline of the tcgen05 MLP kernel (needs multimodal_gar_b200/libmgar_trace.so built with -DMGAR_TC16_TRACE)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
MGAR_LIB=$PWD/multimodal_gar_b200/libmgar_trace.so timeout 300 python bench.py --steps 1 --warmup 1 --quick --no-cpu-baseline > gpurun_out/bq.json 2> gpurun_out/trace.err
grep tc16trace gpurun_out/trace.err | awk '$3!=-1' | tail -32 > gpurun_out/trace.txt; wc -l gpurun_out/trace.txt
