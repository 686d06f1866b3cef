// CUDA-core (SIMT) attention kernels.  These are NOT the product path: they exist so that
//  (a) the tcgen05 kernels can be cross-checked on the GPU at sizes where the CPU oracle is too slow,
//  (b) the S_dmask debug output (return_attn_probs) can be produced from the final softmax_lse and the
//      position-keyed dropout stream, independent of the main kernel's tiling,
//  (c) bring-up has a simple, obviously-correct implementation of every feature of the path.
// One warp owns one query row (forward, dQ) or one key row (dK/dV); fp32 math throughout.
#include "fa_common.cuh"
#include "fa_kernels.h"

namespace fa {

namespace {

constexpr int kWarpsPerCta = 4;
constexpr int kMaxD = 128;

template <bool kBf16>
__device__ __forceinline__ void load_row_to_smem(float* dst, const void* base, int64_t off, int d, int lane) {
  using T = typename Elem<kBf16>::T;
  const T* src = reinterpret_cast<const T*>(base) + off;
  for (int c = lane; c < d; c += 32) dst[c] = Elem<kBf16>::to_float(src[c]);
}

// dot(smem fp32 row, global 16-bit row), 8 elements (16 bytes) per load
template <bool kBf16>
__device__ __forceinline__ float dot_row(const float* a, const void* base, int64_t off, int d) {
  using T = typename Elem<kBf16>::T;
  const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(base) + off);
  float acc = 0.f;
  for (int c = 0; c < d / 8; ++c) {
    const uint4 raw = __ldg(src + c);
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int t = 0; t < 8; ++t) acc = fmaf(a[c * 8 + t], Elem<kBf16>::to_float(e[t]), acc);
  }
  return acc;
}

template <bool kBf16>
__device__ __forceinline__ void load4(float (&out)[4], const void* base, int64_t off) {
  using T = typename Elem<kBf16>::T;
  const uint2 raw = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const T*>(base) + off));
  const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
  for (int t = 0; t < 4; ++t) out[t] = Elem<kBf16>::to_float(e[t]);
}

template <bool kBf16>
__device__ __forceinline__ void store4(void* base, int64_t off, const float (&v)[4]) {
  using T = typename Elem<kBf16>::T;
  T e[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) e[t] = Elem<kBf16>::from_float(v[t]);
  *reinterpret_cast<uint2*>(reinterpret_cast<T*>(base) + off) = *reinterpret_cast<const uint2*>(e);
}

__device__ __forceinline__ float warp_max(float x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
  return x;
}
__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

// ------------------------------------------------------------------------------------------------
// forward: O, softmax_lse and (optionally) S_dmask.  only_sdmask: read lse, write S_dmask only.
// ------------------------------------------------------------------------------------------------
template <bool kBf16>
__global__ void __launch_bounds__(kWarpsPerCta * 32) simt_fwd_kernel(const fa_params p, const BlockMaskView bm,
                                                                    const int only_sdmask) {
  __shared__ float q_s[kWarpsPerCta][kMaxD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bidb = blockIdx.z, bidh = blockIdx.y;
  const SeqInfo seq(p, bidb);
  const int i = blockIdx.x * kWarpsPerCta + warp;   // query row within the sequence
  if (i >= seq.seqlen_q) return;
  const int d = p.d;
  const uint32_t bh = bidb * p.h + bidh;
  const PhiloxKey pk = make_philox_key(p.philox_seed, p.philox_offset);
  const uint32_t c3 = philox_c3(p, pk);
  const float slope = p.alibi_slopes != nullptr ? p.alibi_slopes[bidh] : 0.f;   // ALiBi: z = scale*s - slope*|i - j|
  const bool is_dropout = p.p_dropout > 0.f;
  const uint32_t keep_thr = keep_threshold_u16(p.p_dropout);
  const float rp_dropout = 1.f / (1.f - p.p_dropout);

  load_row_to_smem<kBf16>(q_s[warp], p.q, (int64_t)(seq.q_begin + i) * p.q_row_stride + (int64_t)bidh * p.q_head_stride,
                          d, lane);
  __syncwarp();

  const int k_end = p.is_causal ? min(seq.seqlen_k, i + 1) : seq.seqlen_k;
  const int64_t lse_idx = ((int64_t)bh * p.seqlen_q_rounded) + i;
  float lse;

  if (!only_sdmask) {
    float m = -INFINITY, l = 0.f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const bool own = lane * 4 < d;
    for (int j0 = 0; j0 < k_end; j0 += 32) {
      const int j = j0 + lane;
      const bool valid = j < k_end && bm.on(i, j);
      float s = -INFINITY;   // scaled (and biased) score
      if (valid)
        s = p.softmax_scale * dot_row<kBf16>(q_s[warp], p.k, (int64_t)(seq.k_begin + j) * p.k_row_stride + (int64_t)bidh * p.k_head_stride, d)
            - slope * fabsf((float)(i - j));
      const float m_new = fmaxf(m, warp_max(s));
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = __expf(m - m_use);   // m = -inf -> 0
      float pr = valid ? __expf(s - m_use) : 0.f;
      l = l * alpha + warp_sum(pr);
      m = m_new;
      if (is_dropout && valid) {
        const uint4 r = dropout_rand8(pk, c3, bh, (uint32_t)i, (uint32_t)j >> 3);
        if (rand8_get(r, j & 7) > keep_thr) pr = 0.f;
      }
#pragma unroll
      for (int t = 0; t < 4; ++t) acc[t] *= alpha;
      const int nj = min(32, k_end - j0);
      for (int jj = 0; jj < nj; ++jj) {
        const float pj = __shfl_sync(0xffffffffu, pr, jj);
        if (own && pj != 0.f) {
          float vv[4];
          load4<kBf16>(vv, p.v, (int64_t)(seq.k_begin + j0 + jj) * p.v_row_stride + (int64_t)bidh * p.v_head_stride + lane * 4);
#pragma unroll
          for (int t = 0; t < 4; ++t) acc[t] = fmaf(pj, vv[t], acc[t]);
        }
      }
    }
    const float inv_l = (l == 0.f || l != l) ? 1.f : 1.f / l;
    lse = (l == 0.f || l != l) ? -INFINITY : m + __logf(l);
    if (own) {
      float outv[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) outv[t] = acc[t] * inv_l * rp_dropout;
      store4<kBf16>(p.o, (int64_t)(seq.q_begin + i) * p.o_row_stride + (int64_t)bidh * p.o_head_stride + lane * 4, outv);
    }
    if (lane == 0) p.softmax_lse[lse_idx] = lse;
  } else {
    lse = p.softmax_lse[lse_idx];
  }

  if (p.s_dmask != nullptr) {
    using T = typename Elem<kBf16>::T;
    T* srow = reinterpret_cast<T*>(p.s_dmask) + ((int64_t)bh * p.seqlen_q_rounded + i) * p.seqlen_k_rounded;
    for (int j0 = 0; j0 < seq.seqlen_k; j0 += 32) {
      const int j = j0 + lane;
      if (j >= seq.seqlen_k) break;
      const bool valid = j < k_end && bm.on(i, j);
      float pr = 0.f;
      if (valid) {
        const float s = dot_row<kBf16>(q_s[warp], p.k, (int64_t)(seq.k_begin + j) * p.k_row_stride + (int64_t)bidh * p.k_head_stride, d);
        pr = __expf(p.softmax_scale * s - slope * fabsf((float)(i - j)) - lse);
        if (is_dropout) {
          const uint4 r = dropout_rand8(pk, c3, bh, (uint32_t)i, (uint32_t)j >> 3);
          // sign bit set <=> dropped (src/fmha/softmax.h:232-234); -0.0 still encodes "dropped"
          if (rand8_get(r, j & 7) > keep_thr) pr = -pr;
        }
      }
      srow[j] = Elem<kBf16>::from_float(pr);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// softmax_d = rowsum(dO * O)
// ------------------------------------------------------------------------------------------------
template <bool kBf16>
__global__ void __launch_bounds__(kWarpsPerCta * 32) simt_dot_do_o_kernel(const fa_params p) {
  using T = typename Elem<kBf16>::T;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bidb = blockIdx.z, bidh = blockIdx.y;
  const SeqInfo seq(p, bidb);
  const int i = blockIdx.x * kWarpsPerCta + warp;
  if (i >= p.seqlen_q_rounded) return;
  float acc = 0.f;
  if (i < seq.seqlen_q) {
    const T* o = reinterpret_cast<const T*>(p.o) + (int64_t)(seq.q_begin + i) * p.o_row_stride + (int64_t)bidh * p.o_head_stride;
    const T* g = reinterpret_cast<const T*>(p.dout) + (int64_t)(seq.q_begin + i) * p.do_row_stride + (int64_t)bidh * p.do_head_stride;
    for (int c = lane; c < p.d; c += 32) acc = fmaf(Elem<kBf16>::to_float(o[c]), Elem<kBf16>::to_float(g[c]), acc);
    acc = warp_sum(acc);
  }
  if (lane == 0) p.softmax_d[((int64_t)(bidb * p.h + bidh)) * p.seqlen_q_rounded + i] = acc;
}

// ------------------------------------------------------------------------------------------------
// dQ: one warp per query row
// ------------------------------------------------------------------------------------------------
template <bool kBf16>
__global__ void __launch_bounds__(kWarpsPerCta * 32) simt_bwd_dq_kernel(const fa_params p, const BlockMaskView bm) {
  __shared__ float q_s[kWarpsPerCta][kMaxD];
  __shared__ float do_s[kWarpsPerCta][kMaxD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bidb = blockIdx.z, bidh = blockIdx.y;
  const SeqInfo seq(p, bidb);
  const int i = blockIdx.x * kWarpsPerCta + warp;
  if (i >= seq.seqlen_q) return;
  const int d = p.d;
  const uint32_t bh = bidb * p.h + bidh;
  const PhiloxKey pk = make_philox_key(p.philox_seed, p.philox_offset);
  const uint32_t c3 = philox_c3(p, pk);
  const float slope = p.alibi_slopes != nullptr ? p.alibi_slopes[bidh] : 0.f;   // ALiBi: z = scale*s - slope*|i - j|
  const bool is_dropout = p.p_dropout > 0.f;
  const uint32_t keep_thr = keep_threshold_u16(p.p_dropout);
  const float rp_dropout = 1.f / (1.f - p.p_dropout);

  load_row_to_smem<kBf16>(q_s[warp], p.q, (int64_t)(seq.q_begin + i) * p.q_row_stride + (int64_t)bidh * p.q_head_stride, d, lane);
  load_row_to_smem<kBf16>(do_s[warp], p.dout, (int64_t)(seq.q_begin + i) * p.do_row_stride + (int64_t)bidh * p.do_head_stride, d, lane);
  __syncwarp();
  const int64_t stat_idx = (int64_t)bh * p.seqlen_q_rounded + i;
  const float lse = p.softmax_lse[stat_idx];
  const float D = p.softmax_d[stat_idx];
  const int k_end = p.is_causal ? min(seq.seqlen_k, i + 1) : seq.seqlen_k;
  const bool own = lane * 4 < d;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int j0 = 0; j0 < k_end; j0 += 32) {
    const int j = j0 + lane;
    const bool valid = j < k_end && bm.on(i, j) && lse != -INFINITY;
    float ds = 0.f;
    if (valid) {
      const int64_t koff = (int64_t)(seq.k_begin + j) * p.k_row_stride + (int64_t)bidh * p.k_head_stride;
      const int64_t voff = (int64_t)(seq.k_begin + j) * p.v_row_stride + (int64_t)bidh * p.v_head_stride;
      const float s = dot_row<kBf16>(q_s[warp], p.k, koff, d);
      float dp = dot_row<kBf16>(do_s[warp], p.v, voff, d);
      const float pr = __expf(p.softmax_scale * s - slope * fabsf((float)(i - j)) - lse);
      if (is_dropout) {
        const uint4 r = dropout_rand8(pk, c3, bh, (uint32_t)i, (uint32_t)j >> 3);
        dp = (rand8_get(r, j & 7) > keep_thr) ? 0.f : dp * rp_dropout;
      }
      ds = pr * (dp - D) * p.softmax_scale;
    }
    const int nj = min(32, k_end - j0);
    for (int jj = 0; jj < nj; ++jj) {
      const float dsj = __shfl_sync(0xffffffffu, ds, jj);
      if (own && dsj != 0.f) {
        float kk[4];
        load4<kBf16>(kk, p.k, (int64_t)(seq.k_begin + j0 + jj) * p.k_row_stride + (int64_t)bidh * p.k_head_stride + lane * 4);
#pragma unroll
        for (int t = 0; t < 4; ++t) acc[t] = fmaf(dsj, kk[t], acc[t]);
      }
    }
  }
  if (own) store4<kBf16>(p.dq, (int64_t)(seq.q_begin + i) * p.dq_row_stride + (int64_t)bidh * p.dq_head_stride + lane * 4, acc);
}

// ------------------------------------------------------------------------------------------------
// dK, dV: one warp per key row
// ------------------------------------------------------------------------------------------------
template <bool kBf16>
__global__ void __launch_bounds__(kWarpsPerCta * 32) simt_bwd_dkv_kernel(const fa_params p, const BlockMaskView bm) {
  __shared__ float k_s[kWarpsPerCta][kMaxD];
  __shared__ float v_s[kWarpsPerCta][kMaxD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bidb = blockIdx.z, bidh = blockIdx.y;
  const SeqInfo seq(p, bidb);
  const int j = blockIdx.x * kWarpsPerCta + warp;
  if (j >= seq.seqlen_k) return;
  const int d = p.d;
  const uint32_t bh = bidb * p.h + bidh;
  const PhiloxKey pk = make_philox_key(p.philox_seed, p.philox_offset);
  const uint32_t c3 = philox_c3(p, pk);
  const float slope = p.alibi_slopes != nullptr ? p.alibi_slopes[bidh] : 0.f;   // ALiBi: z = scale*s - slope*|i - j|
  const bool is_dropout = p.p_dropout > 0.f;
  const uint32_t keep_thr = keep_threshold_u16(p.p_dropout);
  const float rp_dropout = 1.f / (1.f - p.p_dropout);

  load_row_to_smem<kBf16>(k_s[warp], p.k, (int64_t)(seq.k_begin + j) * p.k_row_stride + (int64_t)bidh * p.k_head_stride, d, lane);
  load_row_to_smem<kBf16>(v_s[warp], p.v, (int64_t)(seq.k_begin + j) * p.v_row_stride + (int64_t)bidh * p.v_head_stride, d, lane);
  __syncwarp();
  const bool own = lane * 4 < d;
  float dk[4] = {0.f, 0.f, 0.f, 0.f}, dv[4] = {0.f, 0.f, 0.f, 0.f};
  const int i_begin = p.is_causal ? (j / 32) * 32 : 0;
  for (int i0 = i_begin; i0 < seq.seqlen_q; i0 += 32) {
    const int i = i0 + lane;
    bool valid = i < seq.seqlen_q && (!p.is_causal || j <= i) && bm.on(i, j);
    float ds = 0.f, pd = 0.f;
    if (valid) {
      const int64_t stat_idx = (int64_t)bh * p.seqlen_q_rounded + i;
      const float lse = p.softmax_lse[stat_idx];
      if (lse != -INFINITY) {
        const float D = p.softmax_d[stat_idx];
        const int64_t qoff = (int64_t)(seq.q_begin + i) * p.q_row_stride + (int64_t)bidh * p.q_head_stride;
        const int64_t gooff = (int64_t)(seq.q_begin + i) * p.do_row_stride + (int64_t)bidh * p.do_head_stride;
        const float s = dot_row<kBf16>(k_s[warp], p.q, qoff, d);
        float dp = dot_row<kBf16>(v_s[warp], p.dout, gooff, d);
        const float pr = __expf(p.softmax_scale * s - slope * fabsf((float)(i - j)) - lse);
        pd = pr;
        if (is_dropout) {
          const uint4 r = dropout_rand8(pk, c3, bh, (uint32_t)i, (uint32_t)j >> 3);
          const bool keep = rand8_get(r, j & 7) <= keep_thr;
          dp = keep ? dp * rp_dropout : 0.f;
          pd = keep ? pr * rp_dropout : 0.f;
        }
        ds = pr * (dp - D) * p.softmax_scale;
      }
    }
    const int ni = min(32, seq.seqlen_q - i0);
    for (int ii = 0; ii < ni; ++ii) {
      const float dsi = __shfl_sync(0xffffffffu, ds, ii);
      const float pdi = __shfl_sync(0xffffffffu, pd, ii);
      if (own && (dsi != 0.f || pdi != 0.f)) {
        float qq[4], gg[4];
        load4<kBf16>(qq, p.q, (int64_t)(seq.q_begin + i0 + ii) * p.q_row_stride + (int64_t)bidh * p.q_head_stride + lane * 4);
        load4<kBf16>(gg, p.dout, (int64_t)(seq.q_begin + i0 + ii) * p.do_row_stride + (int64_t)bidh * p.do_head_stride + lane * 4);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          dk[t] = fmaf(dsi, qq[t], dk[t]);
          dv[t] = fmaf(pdi, gg[t], dv[t]);
        }
      }
    }
  }
  if (own) {
    store4<kBf16>(p.dk, (int64_t)(seq.k_begin + j) * p.dk_row_stride + (int64_t)bidh * p.dk_head_stride + lane * 4, dk);
    store4<kBf16>(p.dv, (int64_t)(seq.k_begin + j) * p.dv_row_stride + (int64_t)bidh * p.dv_head_stride + lane * 4, dv);
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static inline dim3 row_grid(int rows, const fa_params& p) {
  return dim3((rows + kWarpsPerCta - 1) / kWarpsPerCta, p.h, p.b);
}

cudaError_t launch_simt_fwd(const fa_params& p, const BlockMaskView& bm, bool only_sdmask, cudaStream_t stream) {
  if (p.max_seqlen_q <= 0) return cudaSuccess;
  const dim3 grid = row_grid(p.max_seqlen_q, p);
  if (p.is_bf16) simt_fwd_kernel<true><<<grid, kWarpsPerCta * 32, 0, stream>>>(p, bm, only_sdmask ? 1 : 0);
  else simt_fwd_kernel<false><<<grid, kWarpsPerCta * 32, 0, stream>>>(p, bm, only_sdmask ? 1 : 0); count_launch();
  return cudaGetLastError();
}

cudaError_t launch_dot_do_o(const fa_params& p, cudaStream_t stream) {
  if (p.seqlen_q_rounded <= 0) return cudaSuccess;
  const dim3 grid = row_grid(p.seqlen_q_rounded, p);
  if (p.is_bf16) simt_dot_do_o_kernel<true><<<grid, kWarpsPerCta * 32, 0, stream>>>(p);
  else simt_dot_do_o_kernel<false><<<grid, kWarpsPerCta * 32, 0, stream>>>(p); count_launch();
  return cudaGetLastError();
}

cudaError_t launch_simt_bwd(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream) {
  cudaError_t e = launch_dot_do_o(p, stream);
  if (e != cudaSuccess) return e;
  if (p.max_seqlen_q > 0) {
    const dim3 grid = row_grid(p.max_seqlen_q, p);
    if (p.is_bf16) simt_bwd_dq_kernel<true><<<grid, kWarpsPerCta * 32, 0, stream>>>(p, bm);
    else simt_bwd_dq_kernel<false><<<grid, kWarpsPerCta * 32, 0, stream>>>(p, bm); count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  if (p.max_seqlen_k > 0) {
    const dim3 grid = row_grid(p.max_seqlen_k, p);
    if (p.is_bf16) simt_bwd_dkv_kernel<true><<<grid, kWarpsPerCta * 32, 0, stream>>>(p, bm);
    else simt_bwd_dkv_kernel<false><<<grid, kWarpsPerCta * 32, 0, stream>>>(p, bm); count_launch();
    e = cudaGetLastError();
  }
  return e;
}

}  // namespace fa
