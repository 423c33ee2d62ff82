#!/bin/bash
# tools/build_variant.sh <name> [extra nvcc flags, e.g. -DSPK_REFINE_UNROLL=2]: builds variants/lib_<name>.so for A/B runs (tools/ab_bench.sh)
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p variants/obj_$name
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr $*"
for s in spk_prep spk_hamming spk_mi_exact spk_mi_gemm spk_mi_i8 spk_graph spk_comm spk_noalloc spk_cabi; do
  nvcc $FLAGS -c spydrpick_b200/csrc/$s.cu -o variants/obj_$name/$s.o &
done
wait
nvcc -shared -o variants/lib_$name.so variants/obj_$name/*.o -ldl && rm -rf variants/obj_$name && ls -la variants/lib_$name.so
