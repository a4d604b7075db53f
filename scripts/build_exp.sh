#!/bin/bash
# timing-only experiment builds of the tcgen05 MLP kernel: scripts/build_exp.sh 1 2 3 ... -> multimodal_gar_b200/libmgar_exp<n>.so
# (sa_mlp_tc16.cu compiled with -DMGAR_TC16_EXP=<n>, every other object from the regular build)
cd "$(dirname "$0")/../multimodal_gar_b200/csrc" || exit 1
make > /dev/null || exit 1
for n in "$@"; do
  mkdir -p build_exp
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
       --expt-relaxed-constexpr -DMGAR_TC16_EXP=$n -c -o build_exp/sa_mlp_tc16_$n.o sa_mlp_tc16.cu || exit 1
  objs=$(ls build/*.o | grep -v sa_mlp_tc16.o)
  nvcc -shared -gencode arch=compute_100a,code=sm_100a -o ../libmgar_exp$n.so $objs build_exp/sa_mlp_tc16_$n.o -cudart static || exit 1
  echo built libmgar_exp$n.so
done
