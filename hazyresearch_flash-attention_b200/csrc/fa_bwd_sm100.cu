// FlashAttention backward for sm_100a: TMA -> shared memory -> tcgen05.mma with TMEM accumulators.
// Replaces the native `bwd` / `bwd_block` the reference's Python layer binds (flash_attn/flash_attn_interface.py:31-33,
// flash_attn/flash_blocksparse_attn_interface.py:56-58; native source absent from the reference snapshot).  The math is
// the autograd of attention_ref (tests/test_flash_attn.py:141-155):
//     P = exp(scale*S - lse)   dV = (P o Z / keep)^T dO   dP = (dO V^T) o Z / keep   dS = P o (dP - D)
//     D = rowsum(dO o O)       dQ = scale * dS K          dK = scale * dS^T Q
//
// Three launches:  preprocess (D, -lse*log2e, per row)  ->  main  ->  postprocess (dQ fp32 accumulator -> 16 bit, zeros written
// back: the accumulator is clean for the next call, so a kept workspace never needs a memset).
// Converting dQ INSIDE the main kernel (last CTA of a tile / of a (batch, head) pair, register- and TMA-based, four protocols)
// was built and measured in round 2 and lost at every size, S = 512 included: the conversion's loads queue behind the other
// SMs' dQ reductions, which saturate the L2 atomic path (~170 us per (batch, head) on the converting SM; 0.23 ms of a 2.35 ms
// kernel against 0.10 - 0.16 ms as a separate HBM-speed pass).  Numbers: profiles/r02_dq_signal_ab.log.
//
// Main kernel: one CTA owns a block of 128 keys (K and V tiles stay in shared memory) and streams 128-row Q / dO tiles.
// Everything the softmax-side threads touch is computed TRANSPOSED (keys on TMEM lanes, queries on TMEM columns):
//     S^T  = K Q^T      (SS MMA)      P^T -> TMEM (16 bit, aliases S^T)      dV += P^T dO   (TS MMA, A from TMEM)
//     dP^T = V dO^T     (SS MMA)      dS^T -> shared memory (16 bit)         dK += dS^T Q   (SS MMA)
//                                                                            dQ  = dS K     (SS MMA, both operands MN-major)
// dQ tiles are drained TMEM -> registers -> shared memory and added into an fp32 accumulator in global memory with
// cp.reduce.async.bulk (one 8 KB bulk reduction per 16 head-dim columns); the postprocess kernel scales and converts it.
// TMEM columns: S^T/P^T [0,128) | dP^T, then dQ [128,256) | dV [256,256+DPAD) | dK [384,384+DPAD).
// Warps: 0-7 compute (warpgroup g handles query columns 64g..64g+63 of every tile), 8-11 dQ drain, 12 MMA issuer,
// 13 TMA loader, 14 idle, 15 zero-fill of rows beyond a sequence.
#include <cudaTypedefs.h>

#include <atomic>

#include "fa_common.cuh"
#include "fa_kernels.h"
#include "sm100_ptx.cuh"

// This file is compiled three times (Makefile), each time with another share of the 88 instantiations of the main kernel, so
// that the build is three parallel jobs instead of one two-minute one:
//   FA_BWD_PART 0  host side, preprocess / postprocess kernels, the 128-column main kernels (plain and trimmed)
//   FA_BWD_PART 1  the 64-column main kernels                       -> bwd_dispatch_d64()
//   FA_BWD_PART 2  the ordered-dQ (FA_FLAG_DETERMINISTIC) kernels   -> bwd_dispatch_ordered()
// (-DFA_BWD_FEW_VARIANTS experiment builds are a single part-0 object with two or three instantiations.)
#ifndef FA_BWD_PART
#define FA_BWD_PART 0
#endif

namespace fa {

namespace {

constexpr int kBM = 128;            // queries per tile
constexpr int kBN = 128;            // keys per CTA
constexpr int kBwdThreads = 512;
#ifndef FA_BWD_COMPUTE_REGS      // register split experiments (tools/build_variants.sh); 8*32*compute + 4*32*drain + 4*32*56 <= 65536
#define FA_BWD_COMPUTE_REGS 160
#define FA_BWD_DRAIN_REGS 136
#define FA_BWD_DRAIN_RING 8
#endif
constexpr int kComputeRegs = FA_BWD_COMPUTE_REGS;
constexpr int kDrainRegs = FA_BWD_DRAIN_REGS;
static_assert(8 * 32 * kComputeRegs + 4 * 32 * kDrainRegs + 4 * 32 * 56 <= 65536, "register file");
constexpr int kOtherRegs = 56;
constexpr int kMaxSparseTiles = 512;

template <int DPAD>
struct BwdSmem {
  static constexpr int kTile = 128 * DPAD * 2;            // one [128 x DPAD] 16-bit tile
  static constexpr int kSub = 128 * 128;                  // [128 rows x 64 cols] swizzle-128B sub tile
  static constexpr int kKOff = 0;
  static constexpr int kVOff = kKOff + kTile;
  // Q stages: a Q tile lives from S(t) (issued two slots early) to dK(t), so with two stages the load of Q(t+2) sits on
  // the per-tile critical path (dK(t) done -> TMA -> S(t+2)); d <= 64 has the shared memory for a third stage
  static constexpr int kQStages = DPAD <= 64 ? 3 : 2;
  static constexpr int kQOff = kVOff + kTile;
  static constexpr int kDoOff = kQOff + kQStages * kTile;   // 1 stage
  static constexpr int kDsOff = kDoOff + kTile;           // dS^T: [128 keys x 128 queries] 16 bit
  static constexpr int kDqBufs = 3;                       // dQ staging: per drain warp a ring of 2 KB buffers
  static constexpr int kDqOff = kDsOff + 2 * kSub;
  static constexpr int kStatOff = kDqOff + 4 * kDqBufs * 2048;   // kQStages x { -lse*log2e [128], D [128] } fp32
  static constexpr int kMaskOff = kStatOff + kQStages * 1024;    // dropout keep bits: 2 x [4 key words][128 queries] u32
  static constexpr int kListOff = kMaskOff + 2 * 2048;    // blocksparse: int16 list[512], uint8 flags[512]
  static constexpr int kBarOff = kListOff + kMaxSparseTiles * 3;
  static constexpr int kNumBars = 21;                     // + k_fix, q_fix, do_fix
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kTotal = kTmemPtrOff + 16;
  static constexpr int kDynamicBytes = kTotal + 1024;     // + alignment slack
};
static_assert(BwdSmem<128>::kDynamicBytes <= 232448, "shared memory budget");

}  // namespace

// (named namespace: these two appear in the signatures that cross the three objects)
struct BwdTensorMaps {
  CUtensorMap q, k, v, dout;
};

// Backward scratch, indexed by PACKED row so that it is O(total_q * h * d) whatever the batch looks like: sequence b owns the
// 128-row tile slots  (q_begin >> 7) + b + m,  m = 0 .. ceil(seqlen_q / 128) - 1  (slots of different sequences never overlap,
// and there are at most (total_q >> 7) + b + 1 of them).
struct BwdWs {
  float* nl;         // (h, n_slots * 128): -lse*log2e, -inf for rows that must produce P = 0
  float* dsum;       // (h, n_slots * 128): D = rowsum(dO o O), 0 for invalid rows
  float* dq_accum;   // (h, n_slots, 4 row quarters, DPAD/4, 32, 4) fp32; all zero on entry, and again after the postprocess pass
  // behind dsum (not a member: the kernels that do not use it keep their parameter block): int turn[h][n_slots][4 row quarters],
  // FA_FLAG_DETERMINISTIC -- how many key blocks have added their part to that quarter tile; cleared by every call that uses it
  int n_slots;
};
cudaError_t bwd_dispatch_d64(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws, cudaStream_t stream);
cudaError_t bwd_dispatch_ordered(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws, cudaStream_t stream);

namespace {

__device__ __forceinline__ int tile_slot(int q_begin, int bidb, int m) { return (q_begin >> 7) + bidb + m; }

// Optional timeline tracing of one CTA (build with -DFA_BWD_TRACE; tools/trace_bwd.py): clock64 at pipeline events.
#ifdef FA_BWD_TRACE
__device__ long long g_bwd_trace[64 * 32];
__device__ long long g_bwd_cta[8192 * 12];   // per CTA: smid, globaltimer start, clock start, after setup, compute loop end, epilogue end, globaltimer end
#define FA_CTA_MARK(slot, val)                                                            \
  do {                                                                                    \
    const int cta_lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);   \
    if (cta_lin < 8192) g_bwd_cta[cta_lin * 12 + (slot)] = (val);                          \
  } while (0)
#define FA_TRACE(slot)                                                                                   \
  do {                                                                                                   \
    if (trace_cta && it < 64 && (threadIdx.x & 31) == 0) g_bwd_trace[it * 32 + (slot)] = clock64();      \
  } while (0)
#else
#define FA_TRACE(slot) do { } while (0)
#define FA_CTA_MARK(slot, val) do { } while (0)
#endif
#ifndef FA_POST_TRIM_BATCH
#define FA_POST_TRIM_BATCH 4   // trimmed postprocess pass: 16-column chunks in flight per thread (1 / 2 / 4 measured: profiles/r02_head_dim_trim_ab.log)
#endif

__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];\n" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// dQ finalisation of one tile quarter (32 query rows): fp32 accumulator (L2) -> x scale -> 16 bit -> dq, and zeros written
// back so that the accumulator is clean for the next call.  Quarter layout: [DPAD/16 chunks][4 column groups][32 rows][4].
// `dq_row`: this lane's output row, nullptr when the row lies beyond the sequence.  kBatch chunks are in flight at a time.
template <int DPAD, bool kBf16, int kBatch, bool kTrim>
__device__ __forceinline__ void finalize_dq_quarter(float* acc, uint16_t* dq_row, int d, float scale, int lane) {
  // kTrim: the main kernel ran trimmed and never touched the chunks past ceil16(d) -- they are still zero.  (A template flag, not a
  // run-time count: the predicated form needs 84 registers instead of 70, one resident CTA less per SM, +25 % on the full-width pass.)
  constexpr int kB = kTrim ? FA_POST_TRIM_BATCH : kBatch;        // chunks in flight per thread
  constexpr bool kGuard = kTrim && kB > 2;                       // else: the chunk count is rounded up to whole batches (zeros)
  const int n_exact = kTrim ? (d + 15) >> 4 : DPAD / 16;
  const int n_chunks = kGuard ? n_exact : (n_exact + kB - 1) / kB * kB;
#pragma unroll 1
  for (int k0 = 0; k0 < n_chunks; k0 += kB) {
    float4 v[kB][4];
#pragma unroll
    for (int b = 0; b < kB; ++b)
#pragma unroll
      for (int g = 0; g < 4; ++g)
        v[b][g] = (!kGuard || k0 + b < n_chunks) ? __ldcg(reinterpret_cast<const float4*>(acc + (k0 + b) * 512 + g * 128 + lane * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int b = 0; b < kB; ++b)
#pragma unroll
      for (int g = 0; g < 4; ++g)
        if (!kGuard || k0 + b < n_chunks) __stcg(reinterpret_cast<float4*>(acc + (k0 + b) * 512 + g * 128 + lane * 4), make_float4(0.f, 0.f, 0.f, 0.f));
#pragma unroll
    for (int b = 0; b < kB; ++b) {
      const int col = (k0 + b) * 16;
      uint4 lo, hi;
      lo.x = pack2<kBf16>(v[b][0].x * scale, v[b][0].y * scale);
      lo.y = pack2<kBf16>(v[b][0].z * scale, v[b][0].w * scale);
      lo.z = pack2<kBf16>(v[b][1].x * scale, v[b][1].y * scale);
      lo.w = pack2<kBf16>(v[b][1].z * scale, v[b][1].w * scale);
      hi.x = pack2<kBf16>(v[b][2].x * scale, v[b][2].y * scale);
      hi.y = pack2<kBf16>(v[b][2].z * scale, v[b][2].w * scale);
      hi.z = pack2<kBf16>(v[b][3].x * scale, v[b][3].y * scale);
      hi.w = pack2<kBf16>(v[b][3].z * scale, v[b][3].w * scale);
      if (dq_row != nullptr && col < d) *reinterpret_cast<uint4*>(dq_row + col) = lo;
      if (dq_row != nullptr && col + 8 < d) *reinterpret_cast<uint4*>(dq_row + col + 8) = hi;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// main kernel
// ------------------------------------------------------------------------------------------------
// kMode 1 (trimmed): head dim below DPAD -- issue only the MMA k-steps / output columns / dQ reduction chunks that hold data (a
// separate instantiation: with the counts as run-time values the full-width kernels lose 3 % to the extra registers and branches).
// kMode 2 (deterministic, FA_FLAG_DETERMINISTIC): trimmed, and the key blocks add their dQ parts to a query tile one after the
// other, in key-block order -- dQ becomes bitwise reproducible like dK and dV.  Per (head, tile, 32-row quarter) a counter in the
// workspace holds how many key blocks have added their part; the drain warp of a key block waits for the value that equals its
// rank on that tile, issues its bulk reductions, waits until they have completed (written, not just read), and publishes
// rank + 1.  The ranks follow the order in which the blocks reach a tile anyway (see tile_of below), and a block only ever waits
// for CTAs that are running or done.  A stalled turn traps after 4 s.
template <int DPAD, bool kBf16, bool kCausal, bool kDropout, bool kSparse, bool kAlibi, int kMode>
__global__ void __launch_bounds__(kBwdThreads, 1)
bwd_sm100_kernel(const __grid_constant__ BwdTensorMaps maps, const fa_params p, const BlockMaskView bm, const BwdWs ws,
                 const int n_kblocks, const int bh_group, const PhiloxKey pk) {   // round keys precomputed on the host (constant bank)
  constexpr bool kTrim = kMode != 0;
  constexpr bool kDet = kMode == 2;
  static_assert(!(kDet && kSparse), "deterministic dQ: dense launches only (with a block mask the contributors of a tile are not 0..j)");
  using L = BwdSmem<DPAD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D launch.  causal: longest (lowest key block) first within groups of (batch, head) pairs; else block-major inside
  // each (batch, head) pair
  int bidb, bidh, n_blk;
  if (kCausal) {
    const CtaCoord c = decode_cta(blockIdx.x, n_kblocks, p.b * p.h, bh_group);
    n_blk = kDet ? n_kblocks - 1 - c.x : c.x;   // ordered dQ: SHORTEST first (see the drain warps), so a CTA only waits for earlier ones
    bidb = c.bh / p.h;
    bidh = c.bh - bidb * p.h;
  } else {
    const int pair = blockIdx.x / n_kblocks;
    n_blk = blockIdx.x - pair * n_kblocks;
    bidb = pair / p.h;
    bidh = pair - bidb * p.h;
  }
  const SeqInfo seq(p, bidb);
  const int k0 = n_blk * kBN;
#ifdef FA_BWD_TRACE
  if (threadIdx.x == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    FA_CTA_MARK(0, (long long)smid);
    FA_CTA_MARK(1, (long long)global_timer_ns());
    FA_CTA_MARK(2, clock64());
  }
#endif
  if (k0 >= seq.seqlen_k) return;   // CTA-uniform: no key rows to produce
  const uint32_t bh = bidb * p.h + bidh;
#ifdef FA_BWD_TRACE
  const bool trace_cta = n_blk == 1 && bidh == 0 && bidb == 0;
#endif

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* k_full = bars + 0;
  uint64_t* v_full = bars + 1;
  constexpr int kQ = L::kQStages;
  uint64_t* q_full = bars + 2;    // [kQ <= 3]
  uint64_t* q_empty = bars + 5;   // [kQ <= 3]
  uint64_t* do_full = bars + 8;
  uint64_t* do_empty = bars + 9;
  uint64_t* s_full = bars + 10;
  uint64_t* dp_full = bars + 11;
  uint64_t* p_full = bars + 12;
  uint64_t* ds_full = bars + 13;
  uint64_t* ds_empty = bars + 14;
  uint64_t* dq_full = bars + 15;
  uint64_t* dq_empty = bars + 16;
  uint64_t* dkv_full = bars + 17;
  // landing barriers of tiles that reach past their sequence: the TMA completes on these, warp 15 zeroes the stray rows
  // and only then completes k_full / q_full[s] / do_full for the MMA warp
  uint64_t* k_fix = bars + 18;
  uint64_t* q_fix = bars + 19;
  uint64_t* do_fix = bars + 20;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + L::kTmemPtrOff);
  int16_t* tile_list = reinterpret_cast<int16_t*>(smem + L::kListOff);
  uint8_t* tile_flags = smem + L::kListOff + kMaxSparseTiles * 2;
  __shared__ int s_num_active;

  // Q tiles this CTA visits
  const int n_qt_total = (seq.seqlen_q + kBM - 1) / kBM;
  const int m_start = kCausal ? n_blk : 0;    // top-left aligned causal: queries >= first key of the block
  int n_iter = max(n_qt_total - m_start, 0);

  if (kSparse) {
    const int c = k0 >> 8;
    for (int m = threadIdx.x; m < n_qt_total; m += kBwdThreads) {
      uint32_t f = 0;
      if (c < bm.ncol) {
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const int r = m * 8 + g;
          if (r < bm.nrow && bm.dense[r * bm.ncol + c] != 0) f |= 1u << g;
        }
      }
      tile_flags[m] = static_cast<uint8_t>(f);
    }
    __syncthreads();
    if (warp == 0) {
      int n_act = 0;
      for (int m0 = 0; m0 < n_qt_total; m0 += 32) {
        const int m = m0 + lane;
        const bool f = m < n_qt_total && tile_flags[m] != 0;
        const uint32_t ball = __ballot_sync(0xffffffffu, f);
        if (f) tile_list[n_act + __popc(ball & ((1u << lane) - 1))] = static_cast<int16_t>(m);
        n_act += __popc(ball);
      }
      if (lane == 0) s_num_active = n_act;
    }
    __syncthreads();
    n_iter = s_num_active;
  }
  auto tile_of = [&](int it) -> int { return kSparse ? static_cast<int>(tile_list[it]) : m_start + it; };
  // Isolation of sequences (the reference predicates its loads on row < seqlen and zero-fills): tiles that reach past the
  // sequence hold the next sequence's rows or whatever lies behind the packed batch.  The softmax-side threads mask such
  // entries with selects, but the MMAs would still multiply an exact 0 by a stray NaN: dQ = dS K (stray K rows),
  // dK = dS^T Q and dV = P^T dO (stray rows of the last query tile).  Warp 15 zeroes those rows in shared memory.
  const int k_valid = seq.seqlen_k - k0;                 // valid key rows of this block (>= 1)
  const bool kv_partial = k_valid < kBN;
  const int q_valid_last = seq.seqlen_q - (n_qt_total - 1) * kBM;   // valid rows of the last query tile (1..128)
  const bool q_partial = n_iter > 0 && q_valid_last < kBM && tile_of(n_iter - 1) == n_qt_total - 1;   // then it is the last iteration
  const int slot0 = tile_slot(seq.q_begin, bidb, 0);

  uint8_t* sK = smem + L::kKOff;
  uint8_t* sV = smem + L::kVOff;
  uint8_t* sQ = smem + L::kQOff;
  uint8_t* sDo = smem + L::kDoOff;
  float* sStat = reinterpret_cast<float*>(smem + L::kStatOff);
  if (warp == 13 && elect_one()) {
    // The TMA loader initialises the barriers of its own loads and starts K, V and the first Q / dO tile right away:
    // the loads run under the TMEM allocation and the CTA-wide barrier below instead of after them.
    mbar_init(k_full, 1);
    mbar_init(v_full, 1);
    for (int s = 0; s < kQ; ++s) {
      mbar_init(&q_full[s], 1);
      mbar_init(&q_empty[s], 1);
    }
    mbar_init(do_full, 1);
    mbar_init(do_empty, 1);
    mbar_init(k_fix, 1);
    mbar_init(q_fix, 1);
    mbar_init(do_fix, 1);
    fence_mbar_init();
    tma_prefetch_desc(&maps.q);
    tma_prefetch_desc(&maps.k);
    tma_prefetch_desc(&maps.v);
    tma_prefetch_desc(&maps.dout);
    if (n_iter > 0) {   // nobody waits for K / V otherwise, and a CTA must not exit with loads in flight
      uint64_t* kbar = kv_partial ? k_fix : k_full;
      mbar_arrive_expect_tx(kbar, L::kTile);
#pragma unroll
      for (int c = 0; c < DPAD / 64; ++c) tma_load_3d(sK + c * L::kSub, &maps.k, kbar, c * 64, bidh, seq.k_begin + k0);
      const int m = tile_of(0);
      const bool fix0 = q_partial && n_iter == 1;
      uint64_t* qbar = fix0 ? q_fix : &q_full[0];
      mbar_arrive_expect_tx(qbar, L::kTile + 1024);
#pragma unroll
      for (int c = 0; c < DPAD / 64; ++c)
        tma_load_3d(sQ + c * L::kSub, &maps.q, qbar, c * 64, bidh, seq.q_begin + m * kBM);
      const int64_t stat_off = ((int64_t)bidh * ws.n_slots + slot0 + m) * kBM;
      bulk_load_1d(sStat, ws.nl + stat_off, 512, qbar);
      bulk_load_1d(sStat + 128, ws.dsum + stat_off, 512, qbar);
      mbar_arrive_expect_tx(v_full, L::kTile);
#pragma unroll
      for (int c = 0; c < DPAD / 64; ++c) tma_load_3d(sV + c * L::kSub, &maps.v, v_full, c * 64, bidh, seq.k_begin + k0);
      uint64_t* dobar = fix0 ? do_fix : do_full;
      mbar_arrive_expect_tx(dobar, L::kTile);
#pragma unroll
      for (int c = 0; c < DPAD / 64; ++c)
        tma_load_3d(sDo + c * L::kSub, &maps.dout, dobar, c * 64, bidh, seq.q_begin + m * kBM);
    }
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    mbar_init(s_full, 1);
    mbar_init(dp_full, 1);
    mbar_init(p_full, 256);
    mbar_init(ds_full, 256);
    mbar_init(ds_empty, 1);
    mbar_init(dq_full, 1);
    mbar_init(dq_empty, 128);
    mbar_init(dkv_full, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) FA_CTA_MARK(3, clock64());
  uint8_t* sDs = smem + L::kDsOff;
  uint8_t* sDq = smem + L::kDqOff;
  uint32_t* sMask = reinterpret_cast<uint32_t*>(smem + L::kMaskOff);

  constexpr uint32_t kColS = 0, kColDP = 128, kColDV = 256, kColDK = 384;
  // d <= 64 leaves TMEM columns free: dQ gets its own (448..511) instead of aliasing dP^T, which takes the
  // dQ(t) -> drain -> dP(t+1) hop out of the per-tile dependency chain
  constexpr bool kDqAliased = DPAD > 64;
  constexpr uint32_t kColDQ = kDqAliased ? kColDP : 448;

  if (warp >= 12) {
    setmaxnreg_dec<kOtherRegs>();
    if (warp == 13) {
      // ===================================== TMA loader ========================================
      if (elect_one()) {
        // K, V and tile 0 were issued before the CTA-wide barrier
        for (int it = 1; it < n_iter; ++it) {
          const int m = tile_of(it);
          const int s = it % kQ;
          const bool fix = q_partial && it + 1 == n_iter;
          mbar_wait(&q_empty[s], ((it / kQ) & 1) ^ 1);
          FA_TRACE(20);
          uint64_t* qbar = fix ? q_fix : &q_full[s];
          mbar_arrive_expect_tx(qbar, L::kTile + 1024);
#pragma unroll
          for (int c = 0; c < DPAD / 64; ++c)
            tma_load_3d(sQ + s * L::kTile + c * L::kSub, &maps.q, qbar, c * 64, bidh, seq.q_begin + m * kBM);
          const int64_t stat_off = ((int64_t)bidh * ws.n_slots + slot0 + m) * kBM;
          bulk_load_1d(sStat + s * 256, ws.nl + stat_off, 512, qbar);
          bulk_load_1d(sStat + s * 256 + 128, ws.dsum + stat_off, 512, qbar);
          mbar_wait(do_empty, (it & 1) ^ 1);
          FA_TRACE(21);
          uint64_t* dobar = fix ? do_fix : do_full;
          mbar_arrive_expect_tx(dobar, L::kTile);
#pragma unroll
          for (int c = 0; c < DPAD / 64; ++c)
            tma_load_3d(sDo + c * L::kSub, &maps.dout, dobar, c * 64, bidh, seq.q_begin + m * kBM);
        }
      }
    } else if (warp == 12) {
      // ===================================== MMA issuer ========================================
      // One elected lane runs the whole role.  (Guarding each tcgen05.mma with `lane == 0` inside warp-wide control flow
      // makes the compiler wrap every single instruction in an ELECT / BRA.U.ANY loop -- a dependent chain of ~100 clk
      // per MMA, i.e. the issue rate, not the tensor pipe, bounded the kernel.  Inside an elect.sync region the
      // instructions are emitted back to back from uniform registers.)
      if (n_iter > 0 && elect_one()) {
        constexpr bool leader = true;
        // head dims below DPAD: columns [d, DPAD) of Q, K, V and dO are zeros (TMA fill), so S^T / dP^T only need the k-steps
        // that hold data and dV / dK / dQ only their first ceil16(d) columns; what lies past them in TMEM is never stored
        const int ksteps = kTrim ? (p.d + 15) >> 4 : DPAD / 16;
        const uint32_t idesc_st = make_idesc_f16(kBf16, 128, 128, false, false);            // S^T, dP^T
        const uint32_t idesc_kv = make_idesc_f16(kBf16, 128, ksteps * 16, false, true);     // dV, dK (B = dO / Q, MN-major)
        const uint32_t idesc_dq = make_idesc_f16(kBf16, 128, ksteps * 16, true, true);      // dQ (A = dS^T read MN-major)
        const uint32_t aQ = smem_u32(sQ);
        const uint32_t tS = tmem_base + kColS, tDP = tmem_base + kColDP, tDV = tmem_base + kColDV, tDK = tmem_base + kColDK;
        // descriptor bases (low words; the high word is constant); a byte offset o inside a tile is added as (o >> 4)
        const uint32_t dK_k = make_smem_desc_lo(smem_u32(sK), 16);          // K-major view (SBO = 8 rows)
        const uint32_t dV_k = make_smem_desc_lo(smem_u32(sV), 16);
        const uint32_t dQ_k = make_smem_desc_lo(aQ, 16);
        const uint32_t dDo_k = make_smem_desc_lo(smem_u32(sDo), 16);
        const uint32_t dDs_k = make_smem_desc_lo(smem_u32(sDs), 16);
        const uint32_t dK_mn = make_smem_desc_lo(smem_u32(sK), L::kSub);    // MN-major view (LBO = next 64 columns)
        const uint32_t dQ_mn = make_smem_desc_lo(aQ, L::kSub);
        const uint32_t dDo_mn = make_smem_desc_lo(smem_u32(sDo), L::kSub);
        const uint32_t dDs_mn = make_smem_desc_lo(smem_u32(sDs), L::kSub);
        constexpr uint32_t kQStage = L::kTile >> 4;   // descriptor units between the two Q stages

        // D[keys x queries] = A[keys x d] B[queries x d]^T, both K-major
        auto issue_t = [&](uint32_t d_tmem, uint32_t a_desc, uint32_t b_desc) {
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < DPAD / 16; ++kk) {
              if (kk >= ksteps) break;
              const uint32_t off = ((kk >> 2) * L::kSub + (kk & 3) * 32) >> 4;
              umma_ss_lo(d_tmem, a_desc + off, b_desc + off, idesc_st, kk > 0 ? 1u : 0u);
            }
          }
        };
        // dV[keys x d] += P^T[keys x queries] (TMEM, 16 bit) dO[queries x d]
        auto issue_dv = [&](bool acc) {
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < kBM / 16; ++kk) {
              const uint32_t a_col = (kk < 4) ? kk * 8 : 64 + (kk - 4) * 8;   // warpgroup 1's half lives at column 64
              umma_ts_lo(tDV, tS + a_col, dDo_mn + (kk * 2048 >> 4), idesc_kv, (acc || kk > 0) ? 1u : 0u);
            }
          }
        };
        // dK[keys x d] += dS^T[keys x queries] (smem, K-major) Q[queries x d] (MN-major)
        auto issue_dk = [&](uint32_t q_stage, bool acc) {
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < kBM / 16; ++kk) {
              const uint32_t off = ((kk >> 2) * L::kSub + (kk & 3) * 32) >> 4;
              umma_ss_lo(tDK, dDs_k + off, dQ_mn + (q_stage * kQStage + (kk * 2048 >> 4)), idesc_kv, (acc || kk > 0) ? 1u : 0u);
            }
          }
        };
        // dQ[queries x d] = dS[queries x keys] (= dS^T tile read MN-major) K[keys x d] (MN-major)
        auto issue_dq = [&]() {
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < kBN / 16; ++kk)
              umma_ss_lo(tmem_base + kColDQ, dDs_mn + (kk * 2048 >> 4), dK_mn + (kk * 2048 >> 4), idesc_dq, kk > 0 ? 1u : 0u);
          }
        };
        auto commit = [&](uint64_t* bar) {
          if (leader) umma_commit(bar);
        };
        // non-blocking poll
        auto poll0 = [&](uint64_t* bar, uint32_t parity) -> bool { return mbar_test_wait(bar, parity); };

        mbar_wait(k_full, 0);
        mbar_wait(&q_full[0], 0);
        tc_fence_after();
        issue_t(tS, dK_k, dQ_k);
        commit(s_full);
        mbar_wait(v_full, 0);
        mbar_wait(do_full, 0);
        tc_fence_after();
        issue_t(tDP, dV_k, dDo_k);
        commit(dp_full);
        // software pipeline: slot `it` issues  dQ(it), dK(it) | dV(it+1) | dP(it+1) | S(it+2).  dQ aliases the dP^T
        // columns, so dP(it+1) has to wait for the drain of dQ(it); dV(it+1) and dK(it) fill that gap.
        mbar_wait(p_full, 0);
        tc_fence_after();
        issue_dv(false);
        commit(do_empty);
        if (n_iter > 1) {
          mbar_wait(&q_full[1], 0);
          tc_fence_after();
          issue_t(tS, dK_k, dQ_k + kQStage);
          commit(s_full);
        }
        for (int it = 0; it < n_iter; ++it) {
          const int s = it % kQ;
          mbar_wait(ds_full, it & 1);
          FA_TRACE(0);
          tc_fence_after();
          if constexpr (!kDqAliased) {
            if (it > 0) {   // the dQ(it) columns only have to wait for the drain of dQ(it-1), which finished long ago
              mbar_wait(dq_empty, (it - 1) & 1);
              tc_fence_after();
            }
          }
          issue_dq();
          FA_TRACE(6);
          commit(dq_full);
          FA_TRACE(7);
          issue_dk(s, it > 0);
          commit(&q_empty[s]);
          commit(ds_empty);
          FA_TRACE(1);
          if (it + 1 < n_iter) {
            mbar_wait(do_full, (it + 1) & 1);
            // dV(it+1) needs P(it+1) from the compute warps, dP(it+1) needs dQ(it) drained out of the aliased columns:
            // issue whichever becomes ready first, then the other.
            if constexpr (!kDqAliased) {   // dP(it+1): its columns are free since dS(it) exists; no drain to wait for
              issue_t(tDP, dV_k, dDo_k);
              commit(dp_full);
              FA_TRACE(4);
            }
            bool dv_done = false, dp_done = !kDqAliased;
            while (!(dv_done && dp_done)) {
              if (!dp_done && (dv_done ? (mbar_wait(dq_empty, it & 1), true) : poll0(dq_empty, it & 1))) {
                FA_TRACE(3);
                tc_fence_after();
                issue_t(tDP, dV_k, dDo_k);
                commit(dp_full);
                dp_done = true;
                FA_TRACE(4);
              }
              if (!dv_done && (dp_done ? (mbar_wait(p_full, (it + 1) & 1), true) : poll0(p_full, (it + 1) & 1))) {
                FA_TRACE(2);
                tc_fence_after();
                issue_dv(true);
                dv_done = true;
              }
            }
            commit(do_empty);
            if (it + 2 < n_iter) {
              const int s2 = (it + 2) % kQ;
              mbar_wait(&q_full[s2], ((it + 2) / kQ) & 1);
              tc_fence_after();
              issue_t(tS, dK_k, dQ_k + s2 * kQStage);
              commit(s_full);
              FA_TRACE(5);
            }
          }
        }
        commit(dkv_full);
      }
      __syncwarp();
    } else if (warp == 15) {
      // ===================================== zero-fill of stray rows ===========================
      auto zero_rows = [&](uint32_t tile_addr, int first_row) {   // rows [first_row, 128) of every 64-column sub tile
        const int n16 = (128 - first_row) * 8;                    // 128-byte rows; the swizzle stays inside a row
        for (int idx = lane; idx < n16 * (DPAD / 64); idx += 32) {
          const int sub = idx / n16, r = idx - sub * n16;
          st_shared_v4(tile_addr + sub * L::kSub + first_row * 128 + r * 16, 0u, 0u, 0u, 0u);
        }
        fence_proxy_async();
        __syncwarp();
      };
      if (n_iter > 0 && kv_partial) {
        mbar_wait(k_fix, 0);
        zero_rows(smem_u32(sK), k_valid);
        if (lane == 0) mbar_arrive(k_full);
      }
      if (q_partial) {
        const int s = (n_iter - 1) % kQ;
        mbar_wait(q_fix, 0);
        zero_rows(smem_u32(sQ) + s * L::kTile, q_valid_last);
        if (lane == 0) mbar_arrive(&q_full[s]);
        mbar_wait(do_fix, 0);
        zero_rows(smem_u32(sDo), q_valid_last);
        if (lane == 0) mbar_arrive(do_full);
      }
    }
  } else if (warp >= 8) {
    // ===================================== dQ drain warps ======================================
    // Each warp drains its own 32 query rows: TMEM -> registers (up to kRing 16-column chunks in flight, so the
    // TMEM columns are handed back to the MMA warp long before the last chunk reaches global memory) -> a private
    // ring of 2 KB staging buffers -> one 2 KB cp.reduce.async.bulk per chunk.  No CTA-wide barrier is involved.
    if constexpr (kDrainRegs > 128) setmaxnreg_inc<kDrainRegs>(); else setmaxnreg_dec<kDrainRegs>();   // 128 at launch
    const int wq = warp & 3;
    const uint32_t tDQ = tmem_base + (static_cast<uint32_t>(wq * 32) << 16) + kColDQ;
    uint8_t* stage_gen = sDq + wq * (L::kDqBufs * 2048);
    const uint32_t stage = smem_u32(stage_gen);
    constexpr int kChunks = DPAD / 16;
    constexpr int kRing = kChunks < FA_BWD_DRAIN_RING ? kChunks : FA_BWD_DRAIN_RING;
    const int n_chunks = kTrim ? (p.d + 15) >> 4 : kChunks;   // 16-column chunks that hold data: the others are neither staged nor reduced
    uint32_t n_issued = 0;
    for (int it = 0; it < n_iter; ++it) {
      const int m = tile_of(it);
      float* gdst = ws.dq_accum + ((int64_t)bidh * ws.n_slots + slot0 + m) * (kBM * DPAD) + wq * (32 * DPAD);
      mbar_wait(dq_full, it & 1);
      if (wq == 0) FA_TRACE(16);
      tc_fence_after();
      uint32_t v[kRing][16];
#pragma unroll
      for (int c = 0; c < kRing; ++c) tmem_ld16(tDQ + c * 16, v[c]);
      tmem_wait_ld();
      if (kRing == kChunks) {
        tc_fence_before();
        mbar_arrive(dq_empty);
      }
      int* turn = nullptr;
      int det_rank = 0;
      if constexpr (kDet) {
        // This CTA's rank among the key blocks that add to query tile m.  The turn order should be the order in which the blocks
        // reach the tile anyway, or they end up a hand-over latency apart each.  Causal: block j starts on the diagonal tile
        // m = j and walks down, so tile m sees blocks min(m, last), ..., 1, 0 one tile time apart: rank counts down from the
        // diagonal, and the launch order is SHORTEST block first (decode above), so that block j waits for j + 1, which has
        // a lower blockIdx.  Non-causal: every block walks the tiles in order, rank = j, block j waits for j - 1.
        det_rank = kCausal ? min(m, (seq.seqlen_k + kBN - 1) / kBN - 1) - n_blk : n_blk;
        turn = reinterpret_cast<int*>(ws.dsum + (int64_t)p.h * ws.n_slots * kBM) + (((int64_t)bidh * ws.n_slots + slot0 + m) << 2) + wq;
        if (lane == 0) {
          uint64_t t0 = 0;
          uint32_t spins = 0;
          while (ld_acquire_gpu(turn) != det_rank) {
            __nanosleep(64);
            if ((++spins & 4095u) == 0) {
              const uint64_t now = global_timer_ns();
              if (t0 == 0) t0 = now;
              else if (now - t0 > 4000000000ull) asm volatile("trap;\n");
            }
          }
        }
        fence_proxy_async_global();      // lane 0 issues the reductions: they are ordered behind its acquire
        __syncwarp();
      }
#pragma unroll
      for (int k = 0; k < kChunks; ++k) {
        if (k < n_chunks) {   // kTrim: chunks past the head dim are neither staged nor reduced
          const uint32_t buf = n_issued % L::kDqBufs;
          if (n_issued >= L::kDqBufs) {
            if (lane == 0) bulk_wait_group_read<L::kDqBufs - 1>();   // the bulk op that last read this buffer is done with it
            __syncwarp();
          }
          ++n_issued;
          const uint32_t dst = stage + buf * 2048 + lane * 16;
#ifndef FA_BWD_EXP_NO_DRAIN
#pragma unroll
          for (int g = 0; g < 4; ++g)
            st_shared_v4(dst + g * 512, v[k % kRing][4 * g], v[k % kRing][4 * g + 1], v[k % kRing][4 * g + 2], v[k % kRing][4 * g + 3]);
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            bulk_reduce_add_f32(gdst + k * 512, stage_gen + buf * 2048, 2048);
            bulk_commit_group();
          }
#else
          if (v[k % kRing][0] == 0x7fc12345u) st_shared_v4(dst, 0, 0, 0, 0);   // timing experiment: no staging, no reduction
          if (lane == 0) bulk_commit_group();
#endif
        }
        if (k + kRing < kChunks) {
          tmem_ld16(tDQ + (k + kRing) * 16, v[k % kRing]);
          tmem_wait_ld();
          if (k + kRing == kChunks - 1) {
            tc_fence_before();
            mbar_arrive(dq_empty);
            if (wq == 0) FA_TRACE(17);
          }
        }
      }
      if constexpr (kDet) {
        if (lane == 0) {
          bulk_wait_group<0>();          // this tile's reductions have been performed, not only read out of the staging buffers
          fence_proxy_async_global();
          __threadfence();
          st_release_gpu(turn, det_rank + 1);
        }
        __syncwarp();
      }
      if (wq == 0) FA_TRACE(18);
    }
    if (warp == 8 && lane == 0) FA_CTA_MARK(9, clock64());    // last chunk issued
    if (lane == 0) bulk_wait_group_read<0>();   // the staging buffers may go away; the reductions complete on their own
    if (warp == 8 && lane == 0) FA_CTA_MARK(10, clock64());
  } else {
    // ===================================== compute warps =======================================
    setmaxnreg_inc<kComputeRegs>();
    const int g = warp >> 2;            // query half of the tile
    const int wq = warp & 3;            // TMEM lane quarter
    const int kl = wq * 32 + lane;      // key row within the block (= TMEM lane)
    const int key = k0 + kl;
    const bool key_valid = key < seq.seqlen_k;
    const uint32_t lane_base = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + kColS + g * 64;
    const uint32_t tDP = tmem_base + lane_base + kColDP + g * 64;
    const float scale_log2 = p.softmax_scale * kLog2e;
    const float rp_dropout = 1.f / (1.f - p.p_dropout);
    const uint32_t keep_thr = keep_threshold_u16(p.p_dropout);
    const uint32_t keep_thr_hi = (keep_thr << 16) | 0xFFFFu;
    const uint32_t c3 = kDropout ? philox_c3(p, pk) : 0u;
    const float nslope2 = kAlibi ? -p.alibi_slopes[bidh] * kLog2e : 0.f;   // ALiBi: exponent += -slope*log2e * |i - j|
    const uint32_t ds_row = smem_u32(sDs) + g * L::kSub + kl * 128;

    for (int it = 0; it < n_iter; ++it) {
      const int m = tile_of(it);
      const int s = it % kQ;
      const uint32_t nl_s = smem_u32(sStat + s * 256 + g * 64);
      const uint32_t d_s = smem_u32(sStat + s * 256 + 128 + g * 64);

      uint64_t keepbits = ~0ull;
      if (kDropout) {
        // all 256 compute threads build the tile's keep bits.  One Philox call covers 8 keys of ONE query, while a
        // thread of the transposed tile needs 64 queries of ONE key: a warp takes 32 consecutive queries of the same
        // key group, and a ballot per key turns its 32 answers into the word (key row, query quarter) directly.
        uint32_t* mbuf = sMask + (it & 1) * 512;          // [128 key rows][4 query quarters]
#pragma unroll
        for (int rep = 0; rep < 2; ++rep) {
          const int idx = threadIdx.x + rep * 256;
          const int ql = idx & 127, w = idx >> 7;         // w (key quarter) and ql >> 5 are warp-uniform
          uint32_t mine = 0;
#pragma unroll
          for (int kg = 0; kg < 4; ++kg) {
            const uint4 r = dropout_rand8(pk, c3, bh, static_cast<uint32_t>(m * kBM + ql), static_cast<uint32_t>((k0 >> 3) + w * 4 + kg));
            const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t b_lo = __ballot_sync(0xffffffffu, (rw[e] & 0xFFFFu) <= keep_thr);
              const uint32_t b_hi = __ballot_sync(0xffffffffu, rw[e] <= keep_thr_hi);
              if (lane == kg * 8 + 2 * e) mine = b_lo;
              if (lane == kg * 8 + 2 * e + 1) mine = b_hi;
            }
          }
          mbuf[(w * 32 + lane) * 4 + (ql >> 5)] = mine;
        }
        named_bar_sync(1, 256);
        keepbits = *reinterpret_cast<const uint64_t*>(mbuf + kl * 4 + g * 2);
      }

      // column validity of (this key, query column c): key padding, causal diagonal, blocksparse row groups
      uint64_t colmask = ~0ull;
      bool need_mask = kv_partial;
      if (kv_partial && !key_valid) colmask = 0;
      if (kCausal && m == n_blk) {
        need_mask = true;
        const int first = kl - g * 64;     // query column c is visible iff c >= first
        colmask &= first <= 0 ? ~0ull : (first >= 64 ? 0ull : (~0ull << first));
      }
      if ((m + 1) * kBM > seq.seqlen_q) {  // last, partial query tile: columns beyond the sequence (their Q / dO rows may hold anything)
        need_mask = true;
        const int nvalid = seq.seqlen_q - m * kBM - g * 64;
        colmask &= nvalid >= 64 ? ~0ull : (nvalid <= 0 ? 0ull : ((1ull << nvalid) - 1ull));
      }
      if (kSparse) {
        need_mask = true;
        const uint32_t f = (static_cast<uint32_t>(tile_flags[m]) >> (g * 4)) & 0xFu;
        uint64_t e = 0;
#pragma unroll
        for (int b4 = 0; b4 < 4; ++b4)
          if (f & (1u << b4)) e |= 0xFFFFull << (16 * b4);
        colmask &= e;
      }

      // s_full also covers the tile's -lse*log2e / D vectors: they complete on the same mbarrier as Q, which the MMA
      // warp observed before issuing S^T (TMA -> q_full -> MMA issue -> commit -> s_full is a release/acquire chain)
      mbar_wait(s_full, it & 1);
      if (wq == 0) FA_TRACE(8 + 4 * g);
      tc_fence_after();

      // ---- P^T = exp2(S^T * scale*log2e - lse*log2e); keep fp32 P for dS, write (P o Z) as 16 bit for dV ----
      float pf[64];
      {
        uint32_t(&r)[2][32] = *reinterpret_cast<uint32_t(*)[2][32]>(pf);
        tmem_ld32(tS, r[0]);
        tmem_ld32(tS + 32, r[1]);
        tmem_wait_ld();
      }
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          const float4 nl = lds128(nl_s + (ch * 32 + c4 * 4) * 4);
          const float nlv[4] = {nl.x, nl.y, nl.z, nl.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = ch * 32 + c4 * 4 + e;
            float x = fmaf(pf[c], scale_log2, nlv[e]);
            if (kAlibi) {
              const float rel0 = static_cast<float>(m * kBM + g * 64 - key);   // i - j at column 0 of this half
              // causal: visible entries have i >= j, the bias is linear in the column
              x += kCausal ? fmaf(nslope2, static_cast<float>(c), nslope2 * rel0) : nslope2 * fabsf(rel0 + static_cast<float>(c));
            }
            pf[c] = ex2(x);
          }
        }
        if (need_mask) {
          const uint32_t mb = static_cast<uint32_t>(colmask >> (ch * 32));
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (!((mb >> c) & 1u)) pf[ch * 32 + c] = 0.f;
        }
        uint32_t pp[16];
        if (kDropout) {
          const uint32_t kb = static_cast<uint32_t>(keepbits >> (ch * 32));
#pragma unroll
          for (int c2 = 0; c2 < 16; ++c2) {
            const float lo = ((kb >> (2 * c2)) & 1u) ? pf[ch * 32 + 2 * c2] : 0.f;
            const float hi = ((kb >> (2 * c2 + 1)) & 1u) ? pf[ch * 32 + 2 * c2 + 1] : 0.f;
            pp[c2] = pack2<kBf16>(lo, hi);
          }
        } else {
#pragma unroll
          for (int c2 = 0; c2 < 16; ++c2) pp[c2] = pack2<kBf16>(pf[ch * 32 + 2 * c2], pf[ch * 32 + 2 * c2 + 1]);
        }
        tmem_st16(tS + ch * 16, pp);   // P^T of this half: columns [64g, 64g+32), aliasing the S^T columns already read
      }
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(p_full);
      if (wq == 0) FA_TRACE(9 + 4 * g);

      // ---- dS^T = P^T o (dP^T o Z / keep - D), 16 bit, into swizzled shared memory ----
      mbar_wait(dp_full, it & 1);
      if (wq == 0) FA_TRACE(10 + 4 * g);
      tc_fence_after();
      if (it > 0) mbar_wait(ds_empty, (it - 1) & 1);   // dK / dQ MMAs of the previous tile have read sDs
      // dP^T comes out of TMEM in one go without dropout; the dropout variants, which also hold the keep bits and run out of
      // registers (pf[64] + dpr[64] live at once spilled 70 values per tile in this loop), take it one 32-column half at a
      // time so that the two halves share registers -- one more TMEM round trip per tile in an issue-bound variant
      constexpr bool kSplitDp = kDropout;
      uint32_t dpr[2][32];
      if constexpr (!kSplitDp) {
        tmem_ld32(tDP, dpr[0]);
        tmem_ld32(tDP + 32, dpr[1]);
        tmem_wait_ld();
      }
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        if constexpr (kSplitDp) {
          tmem_ld32(tDP + ch * 32, dpr[ch]);
          tmem_wait_ld();
        }
        if (need_mask) {
          // masked entries: P is exactly 0 there, but dP may be anything (stray rows of a partial tile) and 0 * NaN = NaN
          const uint32_t mb = static_cast<uint32_t>(colmask >> (ch * 32));
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (!((mb >> c) & 1u)) dpr[ch][c] = 0u;
        }
        uint32_t dsp[16];
        const uint32_t kb = static_cast<uint32_t>(keepbits >> (ch * 32));
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          const float4 dd = lds128(d_s + (ch * 32 + c4 * 4) * 4);
          const float dsum[4] = {dd.x, dd.y, dd.z, dd.w};
          float dsv[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = c4 * 4 + e;
            float dp = __uint_as_float(dpr[ch][c]);
            if (kDropout) dp = ((kb >> c) & 1u) ? dp * rp_dropout : 0.f;
            dsv[e] = pf[ch * 32 + c] * (dp - dsum[e]);
          }
          dsp[c4 * 2] = pack2<kBf16>(dsv[0], dsv[1]);
          dsp[c4 * 2 + 1] = pack2<kBf16>(dsv[2], dsv[3]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t chunk = static_cast<uint32_t>((ch * 4 + j) ^ (kl & 7));
          st_shared_v4(ds_row + chunk * 16, dsp[4 * j], dsp[4 * j + 1], dsp[4 * j + 2], dsp[4 * j + 3]);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(ds_full);
      if (wq == 0) FA_TRACE(11 + 4 * g);
    }

    // ---- epilogue: warpgroup 0 stores dV (x 1/keep), warpgroup 1 stores dK (x scale) ----
    if (threadIdx.x == 0) FA_CTA_MARK(4, clock64());
    {
      if (n_iter > 0) {
        mbar_wait(dkv_full, 0);
        if (threadIdx.x == 0) FA_CTA_MARK(7, clock64());
        tc_fence_after();
      }
      const uint32_t tAcc = tmem_base + lane_base + (g == 0 ? kColDV : kColDK);
      const float out_scale = g == 0 ? (kDropout ? rp_dropout : 1.f) : p.softmax_scale;
      // K and V tiles are dead once dkv_full has completed: each warp stages its 32 rows in its slice of them
      const uint32_t stage = smem_u32(g == 0 ? sV : sK) + wq * (32 * DPAD * 2);
      const int row0 = k0 + wq * 32;
      uint16_t* grow0 = g == 0 ? reinterpret_cast<uint16_t*>(p.dv) + (int64_t)(seq.k_begin + row0) * p.dv_row_stride + (int64_t)bidh * p.dv_head_stride
                               : reinterpret_cast<uint16_t*>(p.dk) + (int64_t)(seq.k_begin + row0) * p.dk_row_stride + (int64_t)bidh * p.dk_head_stride;
      store_acc_rows_coalesced<DPAD, kBf16>(tAcc, n_iter > 0, out_scale, stage, grow0, g == 0 ? p.dv_row_stride : p.dk_row_stride,
                                            min(max(seq.seqlen_k - row0, 0), 32), p.d, lane);
    }
    if (threadIdx.x == 0) FA_CTA_MARK(8, clock64());   // dK / dV stored
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
#ifdef FA_BWD_TRACE
  if (threadIdx.x == 0) {
    FA_CTA_MARK(5, clock64());
    FA_CTA_MARK(6, (long long)global_timer_ns());
  }
#endif
}

#if FA_BWD_PART == 0
// ------------------------------------------------------------------------------------------------
// preprocess: D = rowsum(dO o O) -> softmax_d and the workspace copy; -lse*log2e -> workspace (both padded to whole
// 128-row tiles of each sequence).  HBM-bound: 16 threads per query row (16 bytes each), 64 rows per CTA, all eight
// loads of a thread in flight before the first use.  Rows of sequences without keys get dq = 0 here (no key-block CTA
// will ever finalise them).
// ------------------------------------------------------------------------------------------------
constexpr int kPreRows = 64;
template <bool kBf16>
__global__ void __launch_bounds__(256) bwd_preprocess_kernel(const fa_params p, const BwdWs ws, const int n_rowblocks) {
  using T = typename Elem<kBf16>::T;
  const int bidb = blockIdx.x / n_rowblocks, bidh = blockIdx.y;
  const int rb = blockIdx.x - bidb * n_rowblocks;
  const SeqInfo seq(p, bidb);
  const int rows_pad = (seq.seqlen_q + kBM - 1) / kBM * kBM;       // this sequence's tile slots end here
  const int part = threadIdx.x & 15;
  const int row0 = rb * kPreRows + (threadIdx.x >> 4);
  if (rb * kPreRows >= rows_pad && rb * kPreRows >= p.seqlen_q_rounded) return;
  uint4 ov[4], gv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = row0 + i * 16;
    ov[i] = make_uint4(0u, 0u, 0u, 0u);
    gv[i] = ov[i];
    if (row < seq.seqlen_q && part * 8 < p.d) {
      const int64_t r = seq.q_begin + row;
      ov[i] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.o) + r * p.o_row_stride + (int64_t)bidh * p.o_head_stride + part * 8));
      gv[i] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.dout) + r * p.do_row_stride + (int64_t)bidh * p.do_head_stride + part * 8));
    }
  }
  const int64_t bh = (int64_t)bidb * p.h + bidh;
  const int64_t stat0 = ((int64_t)bidh * ws.n_slots + tile_slot(seq.q_begin, bidb, 0)) * kBM;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = row0 + i * 16;
    const bool valid = row < seq.seqlen_q;
    const T* oe = reinterpret_cast<const T*>(&ov[i]);
    const T* ge = reinterpret_cast<const T*>(&gv[i]);
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t < 8; ++t) acc = fmaf(Elem<kBf16>::to_float(oe[t]), Elem<kBf16>::to_float(ge[t]), acc);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (valid && seq.seqlen_k <= 0 && part * 8 < p.d && p.dq != nullptr)
      *reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.dq) + (int64_t)(seq.q_begin + row) * p.dq_row_stride + (int64_t)bidh * p.dq_head_stride + part * 8) =
          make_uint4(0u, 0u, 0u, 0u);
    if (part == 0) {
      if (row < rows_pad) {
        float nl = -INFINITY;
        if (valid) {
          const float lse = p.softmax_lse[bh * p.seqlen_q_rounded + row];
          if (lse != -INFINITY) nl = -lse * kLog2e;
        }
        ws.nl[stat0 + row] = nl;
        ws.dsum[stat0 + row] = valid ? acc : 0.f;
      }
      if (row < p.seqlen_q_rounded) p.softmax_d[bh * p.seqlen_q_rounded + row] = valid ? acc : 0.f;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// postprocess: dq = scale * dq_accum -> 16 bit, caller's strides; zeros written back (accumulator clean for the next call)
// ------------------------------------------------------------------------------------------------
template <bool kBf16, int DPAD, bool kTrim>
__global__ void __launch_bounds__(128) bwd_postprocess_kernel(const fa_params p, const BwdWs ws, const int n_qtiles) {
  const int bidb = blockIdx.x / n_qtiles, bidh = blockIdx.y;
  const int m = blockIdx.x - bidb * n_qtiles;
  const SeqInfo seq(p, bidb);
  if (m * kBM >= seq.seqlen_q) return;
  const int q4 = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = m * kBM + q4 * 32 + lane;
  float* acc = ws.dq_accum + ((int64_t)bidh * ws.n_slots + tile_slot(seq.q_begin, bidb, m)) * (kBM * DPAD) + q4 * (32 * DPAD);
  uint16_t* out = row < seq.seqlen_q ? reinterpret_cast<uint16_t*>(p.dq) + (int64_t)(seq.q_begin + row) * p.dq_row_stride + (int64_t)bidh * p.dq_head_stride
                                     : nullptr;
  finalize_dq_quarter<DPAD, kBf16, 4, kTrim>(acc, out, p.d, p.softmax_scale, lane);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
inline uint64_t align256(uint64_t x) { return (x + 255) & ~uint64_t(255); }

struct BwdWsLayout {
  uint64_t nl_off, dsum_off, accum_off, turn_off, total;
  int n_slots, dpad;
};

BwdWsLayout bwd_ws_layout(const fa_params& p) {
  BwdWsLayout l;
  l.n_slots = (p.total_q >> 7) + p.b + 1;
  l.dpad = p.d <= 64 ? 64 : 128;
  const uint64_t rows = (uint64_t)p.h * l.n_slots * kBM;
  l.accum_off = 0;                                         // [0, nl_off): the "zero on entry / zero on exit" region
  l.nl_off = align256(rows * l.dpad * 4);
  l.dsum_off = l.nl_off + align256(rows * 4);
  l.turn_off = l.dsum_off + rows * 4;                      // deterministic mode's turn counters, right behind dsum (rows * 4 is a multiple of 512)
  l.total = l.turn_off + align256((uint64_t)p.h * l.n_slots * 4 * sizeof(int));
  return l;
}

// The trimmed instantiations exist where they pay: the 128-column tile without dropout and without a block mask (there the
// kernel is bound by the dQ reduction traffic, which shrinks with the head dim; the dropout variants are ALU-bound and the
// 64-column tile gains < 1 %).  The postprocess pass has to agree, hence one predicate for both.
inline bool bwd_runs_trimmed(const fa_params& p, const BlockMaskView& bm) {
  return p.d > 64 && ((p.d + 15) >> 4) < 8 && !(p.p_dropout > 0.f) && bm.dense == nullptr && !(p.flags & FA_FLAG_DETERMINISTIC);
}
#endif   // FA_BWD_PART == 0

template <int DPAD, bool kBf16, bool kCausal, bool kDropout, bool kSparse, bool kAlibi, int kMode = 0>
cudaError_t launch_bwd_variant(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws,
                               cudaStream_t stream) {
  auto kernel = bwd_sm100_kernel<DPAD, kBf16, kCausal, kDropout, kSparse, kAlibi, kMode>;
  constexpr int smem = BwdSmem<DPAD>::kDynamicBytes;
  static std::atomic<uint64_t> configured{0};   // bit per device (as in the forward)
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 64 || !((configured.load(std::memory_order_acquire) >> dev) & 1u)) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    if (dev < 64) configured.fetch_or(uint64_t(1) << dev, std::memory_order_release);
  }
  const int n_kblocks = (p.max_seqlen_k + kBN - 1) / kBN;
  const int64_t n_ctas = (int64_t)n_kblocks * p.h * p.b;
  if (n_ctas > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  const dim3 grid(static_cast<unsigned>(n_ctas), 1, 1);
  const int bh_group = causal_bh_group(p.max_seqlen_q > p.max_seqlen_k ? p.max_seqlen_q : p.max_seqlen_k);
  void* span = timing_begin(stream);
  kernel<<<grid, kBwdThreads, smem, stream>>>(maps, p, bm, ws, n_kblocks, bh_group, make_philox_key(p.philox_seed, p.philox_offset));
  timing_end(span, 1, stream);
  count_launch();
  return cudaGetLastError();
}

// dense launches: causal x dropout x alibi for one tile width, dtype and kernel mode (the trimmed mode has no dropout variants)
template <int DPAD, bool kBf16, int kMode>
cudaError_t dispatch_bwd_dense(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws,
                               cudaStream_t stream) {
  const int sel = (p.is_causal ? 4 : 0) | (p.p_dropout > 0.f ? 2 : 0) | (p.alibi_slopes != nullptr ? 1 : 0);
  if constexpr (kMode != 1) {
    switch (sel) {
      case 2: return launch_bwd_variant<DPAD, kBf16, false, true, false, false, kMode>(maps, p, bm, ws, stream);
      case 3: return launch_bwd_variant<DPAD, kBf16, false, true, false, true, kMode>(maps, p, bm, ws, stream);
      case 6: return launch_bwd_variant<DPAD, kBf16, true, true, false, false, kMode>(maps, p, bm, ws, stream);
      case 7: return launch_bwd_variant<DPAD, kBf16, true, true, false, true, kMode>(maps, p, bm, ws, stream);
      default: break;
    }
  }
  switch (sel & 5) {
    case 0: return launch_bwd_variant<DPAD, kBf16, false, false, false, false, kMode>(maps, p, bm, ws, stream);
    case 1: return launch_bwd_variant<DPAD, kBf16, false, false, false, true, kMode>(maps, p, bm, ws, stream);
    case 4: return launch_bwd_variant<DPAD, kBf16, true, false, false, false, kMode>(maps, p, bm, ws, stream);
    default: return launch_bwd_variant<DPAD, kBf16, true, false, false, true, kMode>(maps, p, bm, ws, stream);
  }
}

// block mask: non-causal, no bias (flash_blocksparse_attn_interface.py:21)
template <int DPAD, bool kBf16>
cudaError_t dispatch_bwd_sparse(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws,
                                cudaStream_t stream) {
  return p.p_dropout > 0.f ? launch_bwd_variant<DPAD, kBf16, false, true, true, false>(maps, p, bm, ws, stream)
                           : launch_bwd_variant<DPAD, kBf16, false, false, true, false>(maps, p, bm, ws, stream);
}

}  // namespace

#if FA_BWD_PART == 1
cudaError_t bwd_dispatch_d64(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws, cudaStream_t stream) {
  if (bm.dense != nullptr)
    return p.is_bf16 ? dispatch_bwd_sparse<64, true>(maps, p, bm, ws, stream) : dispatch_bwd_sparse<64, false>(maps, p, bm, ws, stream);
  return p.is_bf16 ? dispatch_bwd_dense<64, true, 0>(maps, p, bm, ws, stream) : dispatch_bwd_dense<64, false, 0>(maps, p, bm, ws, stream);
}
#elif FA_BWD_PART == 2
cudaError_t bwd_dispatch_ordered(const BwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, const BwdWs& ws, cudaStream_t stream) {
  if (p.d <= 64)
    return p.is_bf16 ? dispatch_bwd_dense<64, true, 2>(maps, p, bm, ws, stream) : dispatch_bwd_dense<64, false, 2>(maps, p, bm, ws, stream);
  return p.is_bf16 ? dispatch_bwd_dense<128, true, 2>(maps, p, bm, ws, stream) : dispatch_bwd_dense<128, false, 2>(maps, p, bm, ws, stream);
}
#else

uint64_t bwd_sm100_workspace_bytes(const fa_params& p) { return bwd_ws_layout(p).total; }
uint64_t bwd_sm100_workspace_zero_bytes(const fa_params& p) { return bwd_ws_layout(p).nl_off; }

#ifdef FA_BWD_TRACE
extern "C" __attribute__((visibility("default"))) int fa_debug_read_bwd_trace(long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_bwd_trace, sizeof(long long) * n);
}
extern "C" __attribute__((visibility("default"))) int fa_debug_read_bwd_cta(long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_bwd_cta, sizeof(long long) * n);
}
#endif

cudaError_t launch_bwd_sm100(const fa_params& p, const BlockMaskView& bm, void* workspace, cudaStream_t stream,
                             const char** err_msg) {
  *err_msg = nullptr;
  const BwdWsLayout l = bwd_ws_layout(p);
  const int n_qtiles = (p.max_seqlen_q + kBM - 1) / kBM;
  if (bm.dense != nullptr && n_qtiles > kMaxSparseTiles) {
    *err_msg = "blocksparse backward supports at most 65536 queries per sequence";
    return cudaErrorInvalidValue;
  }
  if ((int64_t)n_qtiles * p.b > 0x7fffffffLL || p.h > 65535) {
    *err_msg = "backward grid too large (batch x query tiles must fit 31 bits, heads 16 bits)";
    return cudaErrorInvalidValue;
  }
  uint8_t* base = static_cast<uint8_t*>(workspace);
  BwdWs ws;
  ws.nl = reinterpret_cast<float*>(base + l.nl_off);
  ws.dsum = reinterpret_cast<float*>(base + l.dsum_off);
  ws.dq_accum = reinterpret_cast<float*>(base + l.accum_off);
  ws.n_slots = l.n_slots;
  cudaError_t e;
  if (!(p.flags & FA_FLAG_WS_ZEROED)) {
    e = cudaMemsetAsync(base + l.accum_off, 0, l.nl_off - l.accum_off, stream);
    if (e != cudaSuccess) return e;
  }
  if (p.flags & FA_FLAG_DETERMINISTIC) {   // 16 bytes per (head, query tile): a memset node, so that no other kernel changes for it
    e = cudaMemsetAsync(base + l.turn_off, 0, l.total - l.turn_off, stream);
    if (e != cudaSuccess) return e;
  }
  if (n_qtiles > 0 && p.total_q > 0) {
    const int rows_max = max(n_qtiles * kBM, p.seqlen_q_rounded);
    const int n_rowblocks = (rows_max + kPreRows - 1) / kPreRows;
    const dim3 grid(static_cast<unsigned>(n_rowblocks * p.b), p.h, 1);
    if (p.is_bf16) bwd_preprocess_kernel<true><<<grid, 256, 0, stream>>>(p, ws, n_rowblocks);
    else bwd_preprocess_kernel<false><<<grid, 256, 0, stream>>>(p, ws, n_rowblocks);
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  if (p.max_seqlen_k > 0 && p.total_k > 0) {
    BwdTensorMaps maps;
    const int64_t tq = p.total_q > 0 ? p.total_q : 1;
    const void* qptr = p.total_q > 0 ? p.q : p.k;         // never dereferenced when there are no query rows
    const void* doptr = p.total_q > 0 ? p.dout : p.k;
    if (!make_tensor_map_3d(&maps.q, qptr, p.d, p.h, tq, p.q_row_stride, p.q_head_stride, kBM) ||
        !make_tensor_map_3d(&maps.k, p.k, p.d, p.h, p.total_k, p.k_row_stride, p.k_head_stride, kBN) ||
        !make_tensor_map_3d(&maps.v, p.v, p.d, p.h, p.total_k, p.v_row_stride, p.v_head_stride, kBN) ||
        !make_tensor_map_3d(&maps.dout, doptr, p.d, p.h, tq, p.do_row_stride, p.do_head_stride, kBM)) {
      *err_msg = "cuTensorMapEncodeTiled failed (q/k/v/dout must be 16-byte aligned with 16-byte-multiple strides)";
      return cudaErrorInvalidValue;
    }
#ifdef FA_BWD_FEW_VARIANTS   // experiment builds (tools/build_variants.sh): only the headline instantiation, seconds to compile
    if (p.d > 64 && p.is_bf16 && p.is_causal && !(p.p_dropout > 0.f) && bm.dense == nullptr && p.alibi_slopes == nullptr)
      e = launch_bwd_variant<128, true, true, false, false, false>(maps, p, bm, ws, stream);
    else if (p.d > 64 && p.is_bf16 && !p.is_causal && !(p.p_dropout > 0.f) && bm.dense == nullptr && p.alibi_slopes == nullptr)
      e = launch_bwd_variant<128, true, false, false, false, false>(maps, p, bm, ws, stream);
    else if (p.d <= 64 && !p.is_bf16 && p.is_causal && bm.dense == nullptr && p.alibi_slopes == nullptr)
      e = p.p_dropout > 0.f ? launch_bwd_variant<64, false, true, true, false, false>(maps, p, bm, ws, stream)
                            : launch_bwd_variant<64, false, true, false, false, false>(maps, p, bm, ws, stream);
    else e = cudaErrorNotSupported;
#else
    if (p.flags & FA_FLAG_DETERMINISTIC) e = bwd_dispatch_ordered(maps, p, bm, ws, stream);      // dense only (fa_api.cu validates)
    else if (p.d <= 64) e = bwd_dispatch_d64(maps, p, bm, ws, stream);
    else if (bm.dense != nullptr) e = p.is_bf16 ? dispatch_bwd_sparse<128, true>(maps, p, bm, ws, stream) : dispatch_bwd_sparse<128, false>(maps, p, bm, ws, stream);
    else if (bwd_runs_trimmed(p, bm)) e = p.is_bf16 ? dispatch_bwd_dense<128, true, 1>(maps, p, bm, ws, stream) : dispatch_bwd_dense<128, false, 1>(maps, p, bm, ws, stream);
    else e = p.is_bf16 ? dispatch_bwd_dense<128, true, 0>(maps, p, bm, ws, stream) : dispatch_bwd_dense<128, false, 0>(maps, p, bm, ws, stream);
#endif
    if (e != cudaSuccess) return e;
  }
  if (!(p.flags & FA_FLAG_KEEP_DQ_ACCUM) && n_qtiles > 0 && p.total_q > 0) {
    const dim3 grid(static_cast<unsigned>(n_qtiles * p.b), p.h, 1);
    if (l.dpad == 64) {
      if (p.is_bf16) bwd_postprocess_kernel<true, 64, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
      else bwd_postprocess_kernel<false, 64, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
    } else if (bwd_runs_trimmed(p, bm)) {
      if (p.is_bf16) bwd_postprocess_kernel<true, 128, true><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
      else bwd_postprocess_kernel<false, 128, true><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
    } else {
      if (p.is_bf16) bwd_postprocess_kernel<true, 128, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
      else bwd_postprocess_kernel<false, 128, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
    }
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_bwd_dq_finalize(const fa_params& p, void* workspace, cudaStream_t stream) {
  const BwdWsLayout l = bwd_ws_layout(p);
  const int n_qtiles = (p.max_seqlen_q + kBM - 1) / kBM;
  if (n_qtiles <= 0 || p.total_q <= 0) return cudaSuccess;
  if ((int64_t)n_qtiles * p.b > 0x7fffffffLL || p.h > 65535) return cudaErrorInvalidConfiguration;
  BwdWs ws;
  uint8_t* base = static_cast<uint8_t*>(workspace);
  ws.nl = reinterpret_cast<float*>(base + l.nl_off);
  ws.dsum = reinterpret_cast<float*>(base + l.dsum_off);
  ws.dq_accum = reinterpret_cast<float*>(base + l.accum_off);
  ws.n_slots = l.n_slots;
  const dim3 grid(static_cast<unsigned>(n_qtiles * p.b), p.h, 1);
  // every chunk: the launches that accumulated may or may not have run trimmed
  if (l.dpad == 64) {
    if (p.is_bf16) bwd_postprocess_kernel<true, 64, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
    else bwd_postprocess_kernel<false, 64, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
  } else {
    if (p.is_bf16) bwd_postprocess_kernel<true, 128, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
    else bwd_postprocess_kernel<false, 128, false><<<grid, 128, 0, stream>>>(p, ws, n_qtiles);
  }
  count_launch();
  return cudaGetLastError();
}

#endif   // FA_BWD_PART

}  // namespace fa
