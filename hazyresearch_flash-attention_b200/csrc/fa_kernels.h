// Internal launcher declarations shared by the translation units of libfa_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include "fa_common.cuh"

namespace fa {

// every kernel launch of the library bumps this (fa_launch_count() in the C ABI)
void count_launch();

// optional device timing of the main kernels (see fa_timing_enable in the C ABI); kind 0 = forward, 1 = backward
void* timing_begin(cudaStream_t st);
void timing_end(void* handle, int kind, cudaStream_t st);

// --- CUDA-core cross-check kernels (fa_simt.cu) ---
cudaError_t launch_simt_fwd(const fa_params& p, const BlockMaskView& bm, bool only_sdmask, cudaStream_t stream);
cudaError_t launch_simt_bwd(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream);
cudaError_t launch_dot_do_o(const fa_params& p, cudaStream_t stream);

// --- tcgen05 / TMA kernels (fa_fwd_sm100.cu, fa_bwd_sm100.cu) ---
// Return cudaSuccess or the failing CUDA error; *err_msg (static string) describes non-CUDA failures.
cudaError_t launch_fwd_sm100(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream, const char** err_msg);
// `workspace` points at bwd_sm100_workspace_bytes(p) bytes of scratch (256-byte aligned).
cudaError_t launch_bwd_sm100(const fa_params& p, const BlockMaskView& bm, void* workspace, cudaStream_t stream,
                             const char** err_msg);
// (d, h, total) TMA view of a packed (total, h, d) 16-bit tensor: box 64 x 1 x box_rows, 128-byte swizzle (fa_fwd_sm100.cu)
bool make_tensor_map_3d(CUtensorMap* map, const void* base, int d, int h, int64_t total, int64_t row_stride,
                        int64_t head_stride, int box_rows);
// only the dQ conversion pass of the backward (after FA_FLAG_KEEP_DQ_ACCUM calls)
cudaError_t launch_bwd_dq_finalize(const fa_params& p, void* workspace, cudaStream_t stream);
uint64_t bwd_sm100_workspace_bytes(const fa_params& p);
uint64_t bwd_sm100_workspace_zero_bytes(const fa_params& p);   // the scratch starts with this many zero-on-entry / zero-on-exit bytes

// --- small utility kernels (fa_api.cu) ---
cudaError_t launch_expand_blockmask(const int32_t* converted, int nrow, int ncol, uint8_t* dense, cudaStream_t stream);

}  // namespace fa
