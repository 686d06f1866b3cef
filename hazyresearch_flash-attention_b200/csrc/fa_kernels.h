// Internal launcher declarations shared by the translation units of libfa_b200.so.
#pragma once
#include <cuda_runtime.h>

#include "fa_common.cuh"

namespace fa {

// every kernel launch of the library bumps this (fa_launch_count() in the C ABI)
void count_launch();

// --- CUDA-core cross-check kernels (fa_simt.cu) ---
cudaError_t launch_simt_fwd(const fa_params& p, const BlockMaskView& bm, bool only_sdmask, cudaStream_t stream);
cudaError_t launch_simt_bwd(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream);
cudaError_t launch_dot_do_o(const fa_params& p, cudaStream_t stream);

// --- tcgen05 / TMA kernels (fa_fwd_sm100.cu, fa_bwd_sm100.cu) ---
// Return cudaSuccess or the failing CUDA error; *err_msg (static string) describes non-CUDA failures.
cudaError_t launch_fwd_sm100(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream, const char** err_msg);
cudaError_t launch_bwd_sm100(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream, const char** err_msg);
uint64_t bwd_sm100_workspace_bytes(const fa_params& p);

// --- small utility kernels (fa_api.cu) ---
cudaError_t launch_expand_blockmask(const int32_t* converted, int nrow, int ncol, uint8_t* dense, cudaStream_t stream);

}  // namespace fa
