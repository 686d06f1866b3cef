// FlashAttention forward for sm_100a: TMA -> shared-memory ring -> tcgen05.mma with TMEM accumulators,
// warp-specialised (2 softmax warpgroups ping-ponging two 128-row Q tiles, 1 TMA producer warp,
// 1 MMA issuer warp).  Replaces fmha_fprop_loop_kernel / device_1xN_loop of the reference
// (csrc/flash_attn/src/fmha_fprop_kernel_dispatch.cu:32-35, src/fmha_fprop_kernel_1xN.h:226-717) with the
// loop order inverted: a CTA owns 256 query rows and streams K/V blocks, so O and the softmax statistics
// never leave the SM (the reference round-trips an fp32 O partial through global memory per K block).
//
// Per CTA:  S_t = Q_t K_j^T  (SS MMA, 128x128xDPAD)  ->  softmax warpgroup t: online softmax in registers,
//           P_t written back to TMEM as 16-bit (aliasing S_t)  ->  O_t += P_t V_j (TS MMA, 128xDPADx128).
// TMEM columns: S0 [0,128) | S1 [128,256) | O0 [256,256+DPAD) | O1 [384,384+DPAD).
#include <cudaTypedefs.h>

#include <atomic>

#include "fa_common.cuh"
#include "fa_kernels.h"
#include "sm100_ptx.cuh"

namespace fa {

namespace {

constexpr int kBlockM = 128;      // rows per Q tile (= UMMA M)
constexpr int kTilesPerCta = 2;   // Q tiles per CTA (ping-pong)
constexpr int kBlockN = 128;      // keys per K/V block
constexpr int kNumThreads = 384;  // warps 0-3: softmax tile 0, 4-7: softmax tile 1, 8: TMA, 9: MMA, 10-11: idle
constexpr int kSoftmaxRegs = 216;
constexpr int kOtherRegs = 72;
// P is handed to the MMA warp in kPChunks column chunks (32 keys each): the PV MMAs of a chunk are issued while the
// softmax warps are still computing the exponentials of the next ones, which takes most of PV off the per-tile
// dependency chain  S -> softmax -> PV -> next S.
constexpr int kPChunks = 4;
constexpr float kRescaleThreshold = 8.0f;   // lazy rescale: only when the row max grows by > 2^8

template <int DPAD>
struct FwdSmem {
  static constexpr int kStages = DPAD == 128 ? 2 : 4;
  static constexpr int kTileBytes = kBlockM * DPAD * 2;          // one [128 x DPAD] 16-bit tile
  static constexpr int kSubTileBytes = kBlockM * 128;            // [128 rows x 64 cols] swizzle-128B sub tile
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + kTilesPerCta * kTileBytes;
  static constexpr int kVOff = kKOff + kStages * kTileBytes;
  static constexpr int kBarOff = kVOff + kStages * kTileBytes;
  // barriers: q_full[2], s_full[2], o_full[2], p_full[2][kPChunks], k_full[S], k_empty[S], v_full[S], v_empty[S], v_fix
  static constexpr int kNumBars = 7 + 2 * kPChunks + 4 * kStages;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kListOff = kTmemPtrOff + 16;              // blocksparse: active KV block list (int16)
  static constexpr int kMaxListEntries = 1024;
  static constexpr int kTotal = kListOff + kMaxListEntries * 2;
  static constexpr int kDynamicBytes = kTotal + 1024;            // + alignment slack
};

struct FwdTensorMaps {
  CUtensorMap q, k, v;
};

// Optional timeline tracing of one CTA (build with -DFA_BWD_TRACE, i.e. `make trace`; tools/trace_fwd.py)
#ifdef FA_BWD_TRACE
__device__ long long g_fwd_trace[64 * 32];
#define FA_FTRACE(u_, slot)                                                                                \
  do {                                                                                                     \
    if (trace_cta && (u_) < 64 && (threadIdx.x & 31) == 0) g_fwd_trace[(u_) * 32 + (slot)] = clock64();   \
  } while (0)
__device__ long long g_fwd_cta[8192 * 8];   // per CTA: smid, globaltimer start, clock start, after setup, loop end (tile 1), end, globaltimer end, K blocks
#define FA_FCTA(slot, val)                                                                    \
  do {                                                                                        \
    const int cta_lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);       \
    if (cta_lin < 8192) g_fwd_cta[cta_lin * 8 + (slot)] = (val);                              \
  } while (0)
#else
#define FA_FTRACE(u_, slot) do { } while (0)
#define FA_FCTA(slot, val) do { } while (0)
#endif

template <int DPAD, bool kBf16, bool kCausal, bool kDropout, bool kSparse, bool kAlibi>
__global__ void __launch_bounds__(kNumThreads, 1)
fwd_sm100_kernel(const __grid_constant__ FwdTensorMaps maps, const fa_params p, const BlockMaskView bm, const int n_mblocks,
                 const int bh_group, const int pingpong, const PhiloxKey pk) {   // round keys precomputed on the host: they are read straight from the constant bank
  using L = FwdSmem<DPAD>;
  constexpr int ST = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D launch (no 65535 limit on batch or heads).  causal: heaviest (latest) query blocks first within groups of
  // (batch, head) pairs; else block-major inside each (batch, head) pair
  int bidb, bidh, m_block;
  if (kCausal) {
    const CtaCoord c = decode_cta(blockIdx.x, n_mblocks, p.b * p.h, bh_group);
    m_block = n_mblocks - 1 - c.x;
    bidb = c.bh / p.h;
    bidh = c.bh - bidb * p.h;
  } else {
    const int pair = blockIdx.x / n_mblocks;
    m_block = blockIdx.x - pair * n_mblocks;
    bidb = pair / p.h;
    bidh = pair - bidb * p.h;
  }
  const SeqInfo seq(p, bidb);
  const int q0 = m_block * (kBlockM * kTilesPerCta);
  if (q0 >= seq.seqlen_q) return;   // CTA-uniform
#ifdef FA_BWD_TRACE
  const bool trace_cta = bidb == 0 && bidh == 0 && m_block == n_mblocks - 1;
  if (threadIdx.x == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    FA_FCTA(0, (long long)smid);
    FA_FCTA(1, (long long)global_timer_ns());
    FA_FCTA(2, clock64());
  }
#endif

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* q_full = bars + 0;
  uint64_t* s_full = bars + 2;
  uint64_t* o_full = bars + 4;
  uint64_t* p_full = bars + 6;                 // [tile][chunk]
  uint64_t* k_full = p_full + 2 * kPChunks;
  uint64_t* k_empty = k_full + ST;
  uint64_t* v_full = k_empty + ST;
  uint64_t* v_empty = v_full + ST;
  uint64_t* v_fix = v_empty + ST;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + L::kTmemPtrOff);
  int16_t* blk_list = reinterpret_cast<int16_t*>(smem + L::kListOff);
  __shared__ int s_num_active;

  // number of KV blocks this CTA walks, and how many of them each Q tile takes part in (a prefix)
  const int nkv_total = (seq.seqlen_k + kBlockN - 1) / kBlockN;
  int nu = nkv_total;
  int cnt[kTilesPerCta];
#pragma unroll
  for (int t = 0; t < kTilesPerCta; ++t) {
    const int r0 = q0 + t * kBlockM;
    int c = (r0 < seq.seqlen_q) ? nkv_total : 0;
    if (kCausal && c > 0) c = min(c, r0 / kBlockN + 1);
    cnt[t] = c;
  }
  if (kCausal) nu = max(cnt[0], cnt[1]);

  if (kSparse) {
    // Active KV blocks = those with at least one nonzero (16 x 256) mask block among this CTA's rows.
    // flags are written into blk_list[512 + j] first, then compacted in order by warp 0.
    const int r_lo = q0 >> 4;
    const int r_hi = min((min(q0 + kBlockM * kTilesPerCta, seq.seqlen_q) + 15) >> 4, bm.nrow);
    for (int j = threadIdx.x; j < nkv_total; j += kNumThreads) {
      const int c = (j * kBlockN) >> 8;
      bool any = false;
      if (c < bm.ncol)
        for (int r = r_lo; r < r_hi; ++r) any |= bm.dense[r * bm.ncol + c] != 0;
      blk_list[512 + j] = any ? 1 : 0;
    }
    __syncthreads();
    if (warp == 0) {
      int n_act = 0;
      for (int j0 = 0; j0 < nkv_total; j0 += 32) {
        const int j = j0 + lane;
        const bool f = j < nkv_total && blk_list[512 + j] != 0;
        const uint32_t ball = __ballot_sync(0xffffffffu, f);
        if (f) blk_list[n_act + __popc(ball & ((1u << lane) - 1))] = static_cast<int16_t>(j);
        n_act += __popc(ball);
      }
      if (lane == 0) s_num_active = n_act;
    }
    __syncthreads();
    nu = s_num_active;
#pragma unroll
    for (int t = 0; t < kTilesPerCta; ++t) cnt[t] = cnt[t] > 0 ? nu : 0;
  }

  // Isolation of sequences (the reference predicates its loads on row < seqlen and zero-fills): the last key block may
  // reach past the sequence, and its V rows then hold the next sequence's tokens or whatever lies behind the packed
  // batch.  P is exactly 0 there, but 0 * NaN in O += P V would poison valid rows -- so the V tile of such a block lands
  // on its own barrier (v_fix), warp 10 zeroes the rows beyond the sequence and only then completes v_full for the MMA
  // warp.  (Stray K rows only produce S columns that the mask overwrites with -inf; stray Q rows are never stored.)
  const int j_last = nu > 0 ? (kSparse ? static_cast<int>(blk_list[nu - 1]) : nu - 1) : 0;
  const int v_valid_last = seq.seqlen_k - j_last * kBlockN;
  const bool v_partial = nu > 0 && v_valid_last < kBlockN;

  uint8_t* sQ = smem + L::kQOff;
  uint8_t* sK = smem + L::kKOff;
  uint8_t* sV = smem + L::kVOff;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1);
      for (int c = 0; c < kPChunks; ++c) mbar_init(&p_full[t * kPChunks + c], kBlockM);
      mbar_init(&o_full[t], 1);
    }
    fence_mbar_init();
  }
  if (warp == 8 && elect_one()) {
    // The TMA producer initialises the barriers of its own loads and starts Q and the first K / V block right away:
    // the loads run under the TMEM allocation and the CTA-wide barrier below instead of after them.
    mbar_init(&q_full[0], 1);
    mbar_init(&q_full[1], 1);
    mbar_init(v_fix, 1);
    for (int s = 0; s < ST; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    fence_mbar_init();
    tma_prefetch_desc(&maps.q);
    tma_prefetch_desc(&maps.k);
    tma_prefetch_desc(&maps.v);
#pragma unroll
    for (int t = 0; t < kTilesPerCta; ++t) {
      if (cnt[t] > 0) {
        mbar_arrive_expect_tx(&q_full[t], L::kTileBytes);
#pragma unroll
        for (int c = 0; c < DPAD / 64; ++c)
          tma_load_3d(sQ + t * L::kTileBytes + c * L::kSubTileBytes, &maps.q, &q_full[t], c * 64, bidh,
                      seq.q_begin + q0 + t * kBlockM);
      }
    }
    if (nu > 0) {
      const int j0 = kSparse ? blk_list[0] : 0;
      mbar_arrive_expect_tx(&k_full[0], L::kTileBytes);
#pragma unroll
      for (int c = 0; c < DPAD / 64; ++c)
        tma_load_3d(sK + c * L::kSubTileBytes, &maps.k, &k_full[0], c * 64, bidh, seq.k_begin + j0 * kBlockN);
      uint64_t* vbar = (nu == 1 && v_partial) ? v_fix : &v_full[0];
      mbar_arrive_expect_tx(vbar, L::kTileBytes);
#pragma unroll
      for (int c = 0; c < DPAD / 64; ++c)
        tma_load_3d(sV + c * L::kSubTileBytes, &maps.v, vbar, c * 64, bidh, seq.k_begin + j0 * kBlockN);
    }
  }
  __syncwarp();
  if (warp == 0) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) FA_FCTA(3, clock64());


  if (warp >= 8) {
    setmaxnreg_dec<kOtherRegs>();
    if (warp == 8) {
      // ===================================== TMA producer =====================================
      if (elect_one()) {
        for (int u = 1; u < nu; ++u) {   // Q and block 0 were issued before the CTA-wide barrier
          const int j = kSparse ? blk_list[u] : u;
          const int s = u % ST;
          const uint32_t ph = (u / ST) & 1;
          mbar_wait(&k_empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&k_full[s], L::kTileBytes);
#pragma unroll
          for (int c = 0; c < DPAD / 64; ++c)
            tma_load_3d(sK + s * L::kTileBytes + c * L::kSubTileBytes, &maps.k, &k_full[s], c * 64, bidh,
                        seq.k_begin + j * kBlockN);
          mbar_wait(&v_empty[s], ph ^ 1);
          uint64_t* vbar = (u + 1 == nu && v_partial) ? v_fix : &v_full[s];
          mbar_arrive_expect_tx(vbar, L::kTileBytes);
#pragma unroll
          for (int c = 0; c < DPAD / 64; ++c)
            tma_load_3d(sV + s * L::kTileBytes + c * L::kSubTileBytes, &maps.v, vbar, c * 64, bidh,
                        seq.k_begin + j * kBlockN);
        }
      }
    } else if (warp == 10) {
      // ===================================== V tail zero-fill ==================================
      if (v_partial) {
        const int s = (nu - 1) % ST;
        mbar_wait(v_fix, 0);
        const int nrows = kBlockN - v_valid_last;   // v_valid_last >= 1: the block exists
        const uint32_t base = smem_u32(sV) + s * L::kTileBytes + v_valid_last * 128;
        for (int idx = lane; idx < nrows * 8 * (DPAD / 64); idx += 32) {
          const int sub = idx / (nrows * 8), r = idx - sub * (nrows * 8);   // 128-byte rows; the swizzle stays inside a row
          asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};\n" ::"r"(base + sub * L::kSubTileBytes + r * 16), "r"(0u) : "memory");
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&v_full[s]);
      }
    } else if (warp == 9) {
      // ===================================== MMA issuer =======================================
      if (elect_one()) {
        // head dims below DPAD: the columns [d, DPAD) of Q, K and V are zeros (TMA fills what lies beyond the tensor's extent), so
        // S = Q K^T only needs the k-steps that hold data and O += P V only the first ceil16(d) output columns.  The O columns
        // past that are never written by an MMA and never stored (the epilogue stores d columns).
        const int ksteps = (p.d + 15) >> 4;
        const uint32_t idesc_qk = make_idesc_f16(kBf16, kBlockM, kBlockN, false, false);
        const uint32_t idesc_pv = make_idesc_f16(kBf16, kBlockM, ksteps * 16, false, true);
        const uint32_t sQ_addr = smem_u32(sQ), sK_addr = smem_u32(sK), sV_addr = smem_u32(sV);

        // S_t = Q_t K^T : A = Q tile (K-major), B = K tile (K-major); DPAD/16 MMAs of K=16
        auto issue_qk = [&](int t, int stage) {
          const uint32_t a0 = sQ_addr + t * L::kTileBytes, b0 = sK_addr + stage * L::kTileBytes;
#pragma unroll
          for (int kk = 0; kk < DPAD / 16; ++kk) {
            if (kk >= ksteps) break;
            const uint32_t off = (kk >> 2) * L::kSubTileBytes + (kk & 3) * 32;
            umma_ss(tmem_base + t * 128, make_smem_desc_sw128(a0 + off, 16, 1024),
                    make_smem_desc_sw128(b0 + off, 16, 1024), idesc_qk, kk > 0 ? 1u : 0u);
          }
        };
        // O_t += P_t V : A = P_t in TMEM (16-bit packed, 8 columns per K=16), B = V tile (MN-major)
        auto issue_pv = [&](int t, int stage, int chunk, bool accumulate) {
          const uint32_t b0 = sV_addr + stage * L::kTileBytes;
          constexpr int kPer = kBlockN / 16 / kPChunks;
#pragma unroll
          for (int kk = chunk * kPer; kk < (chunk + 1) * kPer; ++kk) {
            umma_ts(tmem_base + 256 + t * 128, tmem_base + t * 128 + kk * 8,
                    make_smem_desc_sw128(b0 + kk * 2048, L::kSubTileBytes, 1024), idesc_pv,
                    (accumulate || kk > 0) ? 1u : 0u);
          }
        };

        if (nu > 0) {
          mbar_wait(&k_full[0], 0);
          tc_fence_after();
#pragma unroll
          for (int t = 0; t < kTilesPerCta; ++t) {
            if (cnt[t] > 0) {
              mbar_wait(&q_full[t], 0);
              tc_fence_after();
              issue_qk(t, 0);
              umma_commit(&s_full[t]);
            }
          }
          umma_commit(&k_empty[0]);
        }
        for (int u = 0; u < nu; ++u) {
          const int s = u % ST;
          const uint32_t ph = (u / ST) & 1;
          const int s_next = (u + 1) % ST;
          const uint32_t ph_next = ((u + 1) / ST) & 1;
          mbar_wait(&v_full[s], ph);
          bool k_next_ready = false;
#pragma unroll
          for (int t = 0; t < kTilesPerCta; ++t) {
            if (u < cnt[t]) {
#pragma unroll
              for (int c = 0; c < kPChunks; ++c) {
                mbar_wait(&p_full[t * kPChunks + c], u & 1);
                if (c == 0) FA_FTRACE(u, 16 + 4 * t);
                tc_fence_after();
                issue_pv(t, s, c, u > 0);
              }
              FA_FTRACE(u, 17 + 4 * t);
              if (u + 1 == cnt[t]) umma_commit(&o_full[t]);
            }
            if (u + 1 < cnt[t]) {
              if (!k_next_ready) {
                mbar_wait(&k_full[s_next], ph_next);
                tc_fence_after();
                k_next_ready = true;
              }
              issue_qk(t, s_next);
              umma_commit(&s_full[t]);
              FA_FTRACE(u, 18 + 4 * t);
            }
          }
          umma_commit(&v_empty[s]);
          if (u + 1 < nu) umma_commit(&k_empty[s_next]);
        }
      }
    }
  } else {
    // ===================================== softmax warpgroups =================================
    setmaxnreg_inc<kSoftmaxRegs>();
    const int t = warp >> 2;
    const int row_in_tile = (warp & 3) * 32 + lane;
    const int qi = q0 + t * kBlockM + row_in_tile;          // query row within the sequence
    const bool row_valid = qi < seq.seqlen_q;
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + t * 128;
    const uint32_t tO = tmem_base + lane_base + 256 + t * 128;
    const float scale_log2 = p.softmax_scale * kLog2e;
    const uint32_t bh = bidb * p.h + bidh;
    const uint32_t keep_thr = keep_threshold_u16(p.p_dropout);
    const uint32_t keep_thr_hi = (keep_thr << 16) | 0xFFFFu;
    const uint32_t c3 = kDropout ? philox_c3(p, pk) : 0u;
    // ALiBi (kAlibi): the scores are moved to the scaled log2 domain with the bias added right after the TMEM load,
    //     x = s * scale*log2e - slope*log2e * |i - j|,
    // and everything downstream works on x with a unit scale.
    const float sl2 = kAlibi ? 1.f : scale_log2;
    const float slope2 = kAlibi ? p.alibi_slopes[bidh] * kLog2e : 0.f;

    float m = -INFINITY, l = 0.f;
    const int my_cnt = cnt[t];
    // Ping-pong of the exponential phases: the two tiles of a CTA share the SM's 16 ex2/clk.  Left alone they drift into
    // lockstep -- both warpgroups load S and take the row max at the same time (MUFU idle), then both exponentiate at
    // half speed.  A turn token (named barriers 2 and 3) lets only one tile at a time into its exponential phase, so the
    // other tile's S MMA, TMEM load and row max run underneath it.  Only over the blocks both tiles walk.
    const int n_turns = pingpong ? min(cnt[0], cnt[1]) : 0;
    if (t == 1 && n_turns > 0) named_bar_arrive(2, 256);      // tile 0 goes first
    for (int u = 0; u < my_cnt; ++u) {
      const int j = kSparse ? blk_list[u] : u;
      mbar_wait(&s_full[t], u & 1);
      if ((warp & 3) == 0) FA_FTRACE(u, 8 * t + 0);
      tc_fence_after();
      float sv[kBlockN];
      {
        uint32_t(&r)[4][32] = *reinterpret_cast<uint32_t(*)[4][32]>(sv);
        tmem_ld32(tS + 0, r[0]);
        tmem_ld32(tS + 32, r[1]);
        tmem_ld32(tS + 64, r[2]);
        tmem_ld32(tS + 96, r[3]);
        tmem_wait_ld();
      }
      if ((warp & 3) == 0) FA_FTRACE(u, 8 * t + 1);
      if (kAlibi) {
        const float rel0 = static_cast<float>(qi - j * kBlockN);   // i - j at column 0
        if (kCausal) {   // visible entries have i >= j: the bias is linear in the column
          const float b0 = -slope2 * rel0;
#pragma unroll
          for (int c = 0; c < kBlockN; ++c) sv[c] = fmaf(sv[c], scale_log2, fmaf(slope2, static_cast<float>(c), b0));
        } else {
#pragma unroll
          for (int c = 0; c < kBlockN; ++c) sv[c] = fmaf(sv[c], scale_log2, -slope2 * fabsf(rel0 - static_cast<float>(c)));
        }
      }
      // ---- masking: key padding, causal (top-left aligned: key <= query), blocksparse ----
      int limit = seq.seqlen_k - j * kBlockN;
      if (kCausal) limit = min(limit, qi + 1 - j * kBlockN);
      if (kSparse) { if (!bm.on(qi, j * kBlockN)) limit = 0; }
      if (limit < kBlockN) {
#pragma unroll
        for (int c = 0; c < kBlockN; ++c)
          if (c >= limit) sv[c] = -INFINITY;
      }
      // ---- row max ----
      float mx0 = sv[0], mx1 = sv[1], mx2 = sv[2], mx3 = sv[3];
#pragma unroll
      for (int c = 4; c < kBlockN; c += 4) {
        mx0 = fmaxf(mx0, sv[c]);
        mx1 = fmaxf(mx1, sv[c + 1]);
        mx2 = fmaxf(mx2, sv[c + 2]);
        mx3 = fmaxf(mx3, sv[c + 3]);
      }
      const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      // ---- lazy rescale of the running output (only when the max grew by more than 2^8) ----
      if (u == 0) {
        m = m_blk;
      } else {
        const bool need = (m_blk - m) * sl2 > kRescaleThreshold;   // false on NaN (-inf - -inf)
        if (__any_sync(0xffffffffu, need)) {
          const float m_new = fmaxf(m, m_blk);
          const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
          const float alpha = ex2((m - m_use) * sl2);
          l *= alpha;
          m = m_new;
#pragma unroll
          for (int c4 = 0; c4 < DPAD / 32; ++c4) {
            uint32_t o[32];
            tmem_ld32(tO + c4 * 32, o);
            tmem_wait_ld();
#pragma unroll
            for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
            tmem_st32(tO + c4 * 32, o);
          }
        }
      }
      // ---- exponentials, row sum, dropout, pack to 16 bit: one 32-key chunk at a time, each handed over as soon as its
      //      tcgen05.st has landed (the wait for chunk c is taken after the arithmetic of chunk c+1) ----
      const float m_use = (m == -INFINITY) ? 0.f : m;
      const float neg_ms = -m_use * sl2;
      if ((warp & 3) == 0) FA_FTRACE(u, 8 * t + 2);
      if (u < n_turns) named_bar_sync(2 + t, 256);            // wait for this tile's turn on the MUFU
      if ((warp & 3) == 0) FA_FTRACE(u, 8 * t + 3);
      float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
      constexpr int kCw = kBlockN / kPChunks;     // keys per chunk
#pragma unroll
      for (int ch = 0; ch < kPChunks; ++ch) {
#pragma unroll
        for (int c = ch * kCw; c < (ch + 1) * kCw; c += 4) {
          ffma2(sv[c], sv[c + 1], sl2, neg_ms);          // packed pairs: half the issue slots of FFMA / FADD
          ffma2(sv[c + 2], sv[c + 3], sl2, neg_ms);
          sv[c] = ex2(sv[c]);
          sv[c + 1] = ex2(sv[c + 1]);
          sv[c + 2] = ex2(sv[c + 2]);
          sv[c + 3] = ex2(sv[c + 3]);
          fadd2(sum0, sum1, sv[c], sv[c + 1]);
          fadd2(sum2, sum3, sv[c + 2], sv[c + 3]);
        }
        if (kDropout) {
#pragma unroll
          for (int g = ch * (kCw / 8); g < (ch + 1) * (kCw / 8); ++g) {
            const uint4 r = dropout_rand8(pk, c3, bh, static_cast<uint32_t>(qi), static_cast<uint32_t>(j * (kBlockN / 8) + g));
            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if ((w[e] & 0xFFFFu) > keep_thr) sv[g * 8 + 2 * e] = 0.f;
              if (w[e] > keep_thr_hi) sv[g * 8 + 2 * e + 1] = 0.f;      // (w >> 16) > thr without the shift
            }
          }
        }
        uint32_t pp[kCw / 2];
#pragma unroll
        for (int c = 0; c < kCw / 2; ++c) pp[c] = pack2<kBf16>(sv[ch * kCw + 2 * c], sv[ch * kCw + 2 * c + 1]);
        if (ch > 0) {
          tmem_wait_st();                       // chunk ch-1 has landed
          tc_fence_before();
          mbar_arrive(&p_full[t * kPChunks + ch - 1]);
        }
        tmem_st16(tS + ch * (kCw / 2), pp);
      }
      if ((warp & 3) == 0) FA_FTRACE(u, 8 * t + 4);
      if (u < n_turns && !(t == 1 && u + 1 == n_turns)) named_bar_arrive(3 - t, 256);   // hand the turn over
      l += (sum0 + sum1) + (sum2 + sum3);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&p_full[t * kPChunks + kPChunks - 1]);
      if ((warp & 3) == 0) FA_FTRACE(u, 8 * t + 5);
    }

    if (threadIdx.x == 128) { FA_FCTA(4, clock64()); FA_FCTA(7, (long long)nu); }
    // ---- epilogue: O = acc / l (x 1/keep), lse = m*scale + ln(l) ----
    if (my_cnt > 0 || (q0 + t * kBlockM) < seq.seqlen_q) {
      const bool dead = (l == 0.f) || (l != l);
      const float inv_l = dead ? 1.f : 1.f / l;
      const float out_scale = inv_l * (kDropout ? 1.f / (1.f - p.p_dropout) : 1.f);
      const float lse = dead ? -INFINITY : (m * sl2 + log2f(l)) * kLn2;
      if (my_cnt > 0) {
        mbar_wait(&o_full[t], 0);
        tc_fence_after();
      }
      // o_full also means every S = QK^T of this tile has completed, so the tile's Q buffer is free: each warp stages
      // its 32 output rows there and writes them back as whole rows (see store_acc_rows_coalesced)
      const int row0 = q0 + t * kBlockM + (warp & 3) * 32;
      const uint32_t stage = smem_u32(sQ) + t * L::kTileBytes + (warp & 3) * (32 * DPAD * 2);
      uint16_t* grow0 = reinterpret_cast<uint16_t*>(p.o) + (int64_t)(seq.q_begin + row0) * p.o_row_stride + (int64_t)bidh * p.o_head_stride;
      store_acc_rows_coalesced<DPAD, kBf16>(tO, my_cnt > 0, out_scale, stage, grow0, p.o_row_stride,
                                            min(max(seq.seqlen_q - row0, 0), 32), p.d, lane);
      if (row_valid) p.softmax_lse[(int64_t)bh * p.seqlen_q_rounded + qi] = lse;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
#ifdef FA_BWD_TRACE
  if (threadIdx.x == 0) {
    FA_FCTA(5, clock64());
    FA_FCTA(6, (long long)global_timer_ns());
  }
#endif
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
PFN_cuTensorMapEncodeTiled_v12000 get_tensor_map_encoder() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
  }
  return fn;
}

}  // namespace

#ifdef FA_BWD_TRACE
extern "C" __attribute__((visibility("default"))) int fa_debug_read_fwd_trace(long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_fwd_trace, sizeof(long long) * n);
}
extern "C" __attribute__((visibility("default"))) int fa_debug_read_fwd_cta(long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_fwd_cta, sizeof(long long) * n);
}
#endif

// (d, h, total) view of a packed (total, h, d) tensor with arbitrary row / head strides; box = 64 x 1 x rows,
// 128-byte swizzle, out-of-bounds elements read as zero (this is what pads head dims to 64 / 128 and
// clips the last rows of the packed batch).
bool make_tensor_map_3d(CUtensorMap* map, const void* base, int d, int h, int64_t total, int64_t row_stride,
                        int64_t head_stride, int box_rows) {
  auto encode = get_tensor_map_encoder();
  if (encode == nullptr) return false;
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(d), static_cast<cuuint64_t>(h), static_cast<cuuint64_t>(total)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(head_stride) * 2, static_cast<cuuint64_t>(row_stride) * 2};
  const cuuint32_t box[3] = {64, 1, static_cast<cuuint32_t>(box_rows)};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

namespace {

template <int DPAD, bool kBf16, bool kCausal, bool kDropout, bool kSparse, bool kAlibi>
cudaError_t launch_fwd_variant(const FwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, cudaStream_t stream) {
  auto kernel = fwd_sm100_kernel<DPAD, kBf16, kCausal, kDropout, kSparse, kAlibi>;
  constexpr int smem = FwdSmem<DPAD>::kDynamicBytes;
  static std::atomic<uint64_t> configured{0};   // bit per device: the attribute call costs ~1.5 us of host time
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 64 || !((configured.load(std::memory_order_acquire) >> dev) & 1u)) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    if (dev < 64) configured.fetch_or(uint64_t(1) << dev, std::memory_order_release);
  }
  const int n_mblocks = (p.max_seqlen_q + kBlockM * kTilesPerCta - 1) / (kBlockM * kTilesPerCta);
  const int64_t n_ctas = (int64_t)n_mblocks * p.h * p.b;
  if (n_ctas > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  const dim3 grid(static_cast<unsigned>(n_ctas), 1, 1);
  const int bh_group = causal_bh_group(p.max_seqlen_q > p.max_seqlen_k ? p.max_seqlen_q : p.max_seqlen_k);
  // exponential-phase ping-pong (see the kernel).  Measured (profiles/r01_kernel_experiments_ab.log, after the FFMA2 change):
  // +7-10 % at d = 64 (+1-7 % with dropout); at d = 128 neutral (causal) to -3 % (non-causal): there the second tile fills
  // the MUFU bubbles the first one leaves around its TMEM stores better than the turn token does.  So: d <= 64 only.
  static const int pp_forced = [] { const char* e = getenv("FA_FWD_PINGPONG"); return e != nullptr ? atoi(e) : -1; }();   // experiments; read once
  const int pingpong = pp_forced >= 0 ? pp_forced : (DPAD == 64 ? 1 : 0);
  void* span = timing_begin(stream);
  kernel<<<grid, kNumThreads, smem, stream>>>(maps, p, bm, n_mblocks, bh_group, pingpong, make_philox_key(p.philox_seed, p.philox_offset));
  timing_end(span, 0, stream);
  count_launch();
  return cudaGetLastError();
}

template <int DPAD, bool kBf16>
cudaError_t dispatch_flags(const FwdTensorMaps& maps, const fa_params& p, const BlockMaskView& bm, cudaStream_t stream) {
  const bool drop = p.p_dropout > 0.f;
  if (bm.dense != nullptr) {
    return drop ? launch_fwd_variant<DPAD, kBf16, false, true, true, false>(maps, p, bm, stream)
                : launch_fwd_variant<DPAD, kBf16, false, false, true, false>(maps, p, bm, stream);
  }
  const int sel = (p.is_causal ? 4 : 0) | (drop ? 2 : 0) | (p.alibi_slopes != nullptr ? 1 : 0);
  switch (sel) {
    case 0: return launch_fwd_variant<DPAD, kBf16, false, false, false, false>(maps, p, bm, stream);
    case 1: return launch_fwd_variant<DPAD, kBf16, false, false, false, true>(maps, p, bm, stream);
    case 2: return launch_fwd_variant<DPAD, kBf16, false, true, false, false>(maps, p, bm, stream);
    case 3: return launch_fwd_variant<DPAD, kBf16, false, true, false, true>(maps, p, bm, stream);
    case 4: return launch_fwd_variant<DPAD, kBf16, true, false, false, false>(maps, p, bm, stream);
    case 5: return launch_fwd_variant<DPAD, kBf16, true, false, false, true>(maps, p, bm, stream);
    case 6: return launch_fwd_variant<DPAD, kBf16, true, true, false, false>(maps, p, bm, stream);
    default: return launch_fwd_variant<DPAD, kBf16, true, true, false, true>(maps, p, bm, stream);
  }
}

}  // namespace

// Head-dim dispatch mirrors run_fmha_fprop (src/fmha_fprop_kernel_dispatch.cu:90-134): d <= 64 -> 64, else 128.
cudaError_t launch_fwd_sm100(const fa_params& p, const BlockMaskView& bm, cudaStream_t stream, const char** err_msg) {
  *err_msg = nullptr;
  if (p.max_seqlen_q <= 0 || p.total_q <= 0) return cudaSuccess;
  if (bm.dense != nullptr && (p.max_seqlen_k + kBlockN - 1) / kBlockN > 512) {
    *err_msg = "blocksparse forward supports at most 65536 keys per sequence";
    return cudaErrorInvalidValue;
  }
  FwdTensorMaps maps;
  // K/V tensors may be empty (total_k == 0): TMA needs a nonzero extent, the kernel then never issues a load.
  const int64_t tk = p.total_k > 0 ? p.total_k : 1;
  if (!make_tensor_map_3d(&maps.q, p.q, p.d, p.h, p.total_q, p.q_row_stride, p.q_head_stride, kBlockM) ||
      !make_tensor_map_3d(&maps.k, p.k, p.d, p.h, tk, p.k_row_stride, p.k_head_stride, kBlockN) ||
      !make_tensor_map_3d(&maps.v, p.v, p.d, p.h, tk, p.v_row_stride, p.v_head_stride, kBlockN)) {
    *err_msg = "cuTensorMapEncodeTiled failed (q/k/v must be 16-byte aligned with 16-byte-multiple strides)";
    return cudaErrorInvalidValue;
  }
  if (p.d <= 64) {
    return p.is_bf16 ? dispatch_flags<64, true>(maps, p, bm, stream) : dispatch_flags<64, false>(maps, p, bm, stream);
  }
  return p.is_bf16 ? dispatch_flags<128, true>(maps, p, bm, stream) : dispatch_flags<128, false>(maps, p, bm, stream);
}

}  // namespace fa
