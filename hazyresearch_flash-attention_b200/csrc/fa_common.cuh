// Shared host/device definitions for the attention kernels: the kernel argument block, per-sequence
// bounds, the position-keyed Philox dropout stream and small numeric helpers.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/fa_b200.h"

namespace fa {

#define FA_HD __host__ __device__ __forceinline__

// ------------------------------------------------------------------------------------------------
// Philox4x32, 7 rounds (same round count as the reference RNG, csrc/flash_attn/src/philox.cuh:30-59).
// The counter is keyed by POSITION, not by thread layout:
//     counter = (query row in sequence, key index / 8, batch * h + head, offset low 32 bits)
//     key     = (seed low 32 bits, seed high 32 bits ^ offset high 32 bits)
// One call yields 8 x 16-bit uniforms for 8 consecutive keys; a position is kept iff
// u16 <= floor((1 - p) * 65535)  (16-bit compare as in src/fmha/softmax.h:268-279, fmha_api.cpp:104).
// Forward, backward and the S_dmask debug kernel therefore agree for any tiling.
// ------------------------------------------------------------------------------------------------
struct PhiloxKey {
  uint32_t rk0[7], rk1[7];   // the seven round keys (key + r * Weyl constant): CTA-uniform, computed once, so the
  uint32_t c3;               // per-call cost is 2 wide multiplies + 2 three-input XORs per round and nothing else
};

FA_HD PhiloxKey make_philox_key(uint64_t seed, uint64_t offset) {
  PhiloxKey k;
  const uint32_t k0 = static_cast<uint32_t>(seed);
  const uint32_t k1 = static_cast<uint32_t>(seed >> 32) ^ static_cast<uint32_t>(offset >> 32);
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    k.rk0[r] = k0 + static_cast<uint32_t>(r) * 0x9E3779B9u;
    k.rk1[r] = k1 + static_cast<uint32_t>(r) * 0xBB67AE85u;
  }
  k.c3 = static_cast<uint32_t>(offset);
  return k;
}

FA_HD uint4 philox4x32_7(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const PhiloxKey& key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    const uint64_t p0 = static_cast<uint64_t>(M0) * c0, p1 = static_cast<uint64_t>(M1) * c2;
    const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ key.rk0[r];
    const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ key.rk1[r];
    c1 = static_cast<uint32_t>(p1);
    c3 = static_cast<uint32_t>(p0);
    c0 = n0;
    c2 = n2;
  }
  return make_uint4(c0, c1, c2, c3);
}

// 8 uniforms (16 bit each) for keys 8*kgroup .. 8*kgroup+7 of query row qrow.
FA_HD uint4 dropout_rand8(const PhiloxKey& key, uint32_t c3, uint32_t bh, uint32_t qrow, uint32_t kgroup) {
  return philox4x32_7(qrow, kgroup, bh, c3, key);
}
FA_HD uint32_t rand8_get(const uint4& r, int e) {   // e in [0, 8)
  const uint32_t w = (e >> 1) == 0 ? r.x : (e >> 1) == 1 ? r.y : (e >> 1) == 2 ? r.z : r.w;
  return (e & 1) ? (w >> 16) : (w & 0xFFFFu);
}
FA_HD uint32_t keep_threshold_u16(float p_dropout) {
  const float keep = 1.f - p_dropout;
  return static_cast<uint32_t>(floorf(keep * 65535.0f));
}

// ------------------------------------------------------------------------------------------------
// Per-sequence bounds (restates BlockInfoPadded, csrc/flash_attn/src/fmha_kernel.h:45-76).
// ------------------------------------------------------------------------------------------------
struct SeqInfo {
  int q_begin, k_begin, seqlen_q, seqlen_k;
  // seqused_* (optional): the sequence uses only that many rows from its begin; cu_seqlens[b + 1] is then not read
  __device__ __forceinline__ SeqInfo(const fa_params& p, int b) {
    q_begin = p.cu_seqlens_q[b];
    k_begin = p.cu_seqlens_k[b];
    seqlen_q = p.seqused_q != nullptr ? p.seqused_q[b] : p.cu_seqlens_q[b + 1] - q_begin;
    seqlen_k = p.seqused_k != nullptr ? p.seqused_k[b] : p.cu_seqlens_k[b + 1] - k_begin;
  }
};

// Low word of the Philox offset a kernel uses: the host value plus the optional device-side counter
// (fa_params.philox_offset_ptr -- graph-capturable dropout).  Only counter word c3 depends on it.
__device__ __forceinline__ uint32_t philox_c3(const fa_params& p, const PhiloxKey& pk) {
  return p.philox_offset_ptr != nullptr ? pk.c3 + static_cast<uint32_t>(*p.philox_offset_ptr) : pk.c3;
}

template <bool kBf16>
struct Elem;
template <>
struct Elem<true> {
  using T = __nv_bfloat16;
  static __device__ __forceinline__ float to_float(T x) { return __bfloat162float(x); }
  static __device__ __forceinline__ T from_float(float x) { return __float2bfloat16_rn(x); }
};
template <>
struct Elem<false> {
  using T = __half;
  static __device__ __forceinline__ float to_float(T x) { return __half2float(x); }
  static __device__ __forceinline__ T from_float(float x) { return __float2half_rn(x); }
};

// ------------------------------------------------------------------------------------------------
// CTA order for causal launches.  Work per CTA varies (block x of a sequence has x+1 or nx-x units), so the launch is
// a 1-D grid ordered "longest first" inside groups of `bh_group` (batch, head) pairs: the tail of the grid is made of
// the lightest CTAs, while CTAs that run concurrently still come from few (batch, head) pairs and share K/V (forward)
// or Q/dO and the dQ accumulator (backward) in L2.
// ------------------------------------------------------------------------------------------------
constexpr int kBhGroupMax = 24;
struct CtaCoord {
  int x, bh;   // x = rank by weight inside the sequence (0 = heaviest), bh = batch * h + head
};
__device__ __forceinline__ CtaCoord decode_cta(int lin, int nx, int n_bh, int bh_group) {
  const int per_group = bh_group * nx;
  const int group = lin / per_group;
  const int r = lin - group * per_group;
  const int gsize = min(bh_group, n_bh - group * bh_group);
  CtaCoord c;
  c.x = r / gsize;
  c.bh = group * bh_group + (r - c.x * gsize);
  return c;
}
// Group size, measured on B200 at d = 128 (b x S = 64K tokens, h = 12; profiles/r01_bh_group_sweep.log): the best group
// holds about 64K tokens' worth of (batch, head) pairs -- 16 pairs at S = 4096, 6-8 at 8192, 2-4 at 16384, 24-32 at
// 2048 -- i.e. roughly 64 MB of Q + dO + fp32 dQ accumulator in the backward, half the L2.  Larger groups thrash the L2
// (-20 % at one group for the whole batch), much smaller ones lengthen the tail.  FA_BH_GROUP overrides it (experiments).
inline int causal_bh_group(int max_seqlen) {
  static const int forced = [] { const char* e = getenv("FA_BH_GROUP"); return e != nullptr ? atoi(e) : 0; }();   // read once
  if (forced > 0) return forced;
  const int g = 65536 / (max_seqlen > 0 ? max_seqlen : 1);
  return g < 1 ? 1 : (g > kBhGroupMax ? kBhGroupMax : g);
}

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// Blocksparse helper: the converted mask (ncol, nrow) lists, for each 256-wide key block column, the
// 16-row query blocks that attend to it (x4, low bits = first/last flags), padded with -1
// (flash_attn/flash_blocksparse_attn_interface.py:8-40).  The kernels work from the equivalent dense
// (nrow, ncol) 0/1 byte mask, rebuilt on device by expand_blockmask_kernel.
struct BlockMaskView {
  const uint8_t* dense;   // (nrow, ncol) or nullptr
  int nrow, ncol;
  __device__ __forceinline__ bool on(int qrow, int kcol) const {
    if (dense == nullptr) return true;
    const int r = qrow >> 4, c = kcol >> 8;
    if (r >= nrow || c >= ncol) return false;
    return dense[r * ncol + c] != 0;
  }
};

}  // namespace fa
