#!/bin/bash
# Experiment builds of the backward (A/B timing on the GPU box): csrc/exp/libfa_<tag>.so, each with only the headline
# instantiation of bwd_sm100_kernel (-DFA_BWD_FEW_VARIANTS) plus extra -D flags.  Usage: tools/build_variants.sh tag "-DX=1 ..." [tag flags]...
# Select one at run time with FA_B200_LIB=hazyresearch_flash-attention_b200/csrc/exp/libfa_<tag>.so (debug override of flash_attn_cuda.py).
set -e
cd "$(dirname "$0")/../hazyresearch_flash-attention_b200/csrc"
mkdir -p exp
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++17 $ARCH -lineinfo --use_fast_math --expt-relaxed-constexpr -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
while [ $# -ge 2 ]; do
  tag=$1; extra=$2; shift 2
  ( $NVCC $FLAGS -DFA_BWD_FEW_VARIANTS $extra -Xptxas -v -c fa_bwd_sm100.cu -o exp/bwd_$tag.o 2> exp/bwd_$tag.ptxas.log || { cat exp/bwd_$tag.ptxas.log; exit 1; }
    $NVCC $ARCH -shared -o exp/libfa_$tag.so build/fa_api.o build/fa_simt.o build/fa_fwd_sm100.o exp/bwd_$tag.o -cudart static ) &
done
wait
ls -la exp/*.so
