#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bq.json 2> gpurun_out/bench.err
grep tctrace gpurun_out/bench.err | awk '$3!=-1' | head -32 > gpurun_out/trace.txt; wc -l gpurun_out/trace.txt
