// FlashAttention backward for sm_100a (tcgen05 / TMA).  INTERIM: until the tcgen05 backward lands the AUTO
// path runs the CUDA-core kernels of fa_simt.cu (still a GPU path inside this library -- never a CPU or
// library fallback).
#include "fa_common.cuh"
#include "fa_kernels.h"

namespace fa {

uint64_t bwd_sm100_workspace_bytes(const fa_params& p) { (void)p; return 0; }

cudaError_t launch_bwd_sm100(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream, const char** err_msg) {
  *err_msg = nullptr;
  return launch_simt_bwd(p, bm, stream);
}

}  // namespace fa
