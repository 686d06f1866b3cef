// Raw sm_100a PTX wrappers used by the attention kernels: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (alloc / mma / commit / ld / st / fences) and the UMMA shared-memory / instruction
// descriptors.  No CUTLASS / CuTe: everything here is inline PTX written for this project.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace fa {

#define FA_DEVICE __device__ __forceinline__

FA_DEVICE uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

FA_DEVICE bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ------------------------------------------------------------------------------------------------
// mbarrier
// ------------------------------------------------------------------------------------------------
FA_DEVICE void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
FA_DEVICE void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
FA_DEVICE void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
// generic <-> async proxy ordering on global memory (the deterministic dQ turns: bulk reductions of two CTAs on the same lines)
FA_DEVICE void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;\n" ::: "memory"); }

FA_DEVICE void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
FA_DEVICE void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
FA_DEVICE bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// non-blocking probe (try_wait may suspend the thread for a hardware-defined time; test_wait never does)
FA_DEVICE bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// A wait that has not completed after FA_WAIT_TIMEOUT_NS traps instead of hanging the GPU: a protocol bug then
// surfaces as a launch failure the host can report.  The timer is only read after a failed poll.
#ifndef FA_WAIT_TIMEOUT_NS
#define FA_WAIT_TIMEOUT_NS 4000000000ull
#endif
FA_DEVICE uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
FA_DEVICE void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  // try_wait suspends the thread in hardware for a bounded time; the (slow) global timer is only consulted every
  // 256 failed polls so that it never sits between an arrival and the wake-up.
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 255u) == 0) {
      const uint64_t now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > FA_WAIT_TIMEOUT_NS) { asm volatile("trap;\n"); }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// TMA
// ------------------------------------------------------------------------------------------------
FA_DEVICE void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];\n" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 3-D tiled load: coordinates (c0 = innermost/head-dim element, c1 = head, c2 = token row)
FA_DEVICE void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];\n"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

// 1-D bulk copy global -> shared, completion on an mbarrier (size multiple of 16, 16-byte aligned both sides)
FA_DEVICE void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// 1-D bulk reduction shared -> global: gdst[i] += ssrc[i] (fp32), tracked by the thread's bulk async-group
FA_DEVICE void bulk_reduce_add_f32(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile(
      "cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;\n"
      ::"l"(reinterpret_cast<uint64_t>(gdst)), "r"(smem_u32(smem_src)), "r"(bytes)
      : "memory");
}
// 16-byte vector reduction into global memory (REDG.E.ADD.F32x4): no return value, registers are free at issue.
FA_DEVICE void red_add_v4_f32(float* gptr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n"
               :: "l"(gptr), "f"(__uint_as_float(a)), "f"(__uint_as_float(b)), "f"(__uint_as_float(c)), "f"(__uint_as_float(d))
               : "memory");
}
FA_DEVICE int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.b32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
FA_DEVICE void st_release_gpu(int* p, int v) { asm volatile("st.release.gpu.global.b32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory"); }
FA_DEVICE void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
template <int N>
FA_DEVICE void bulk_wait_group_read() { asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(N) : "memory"); }
template <int N>
FA_DEVICE void bulk_wait_group() { asm volatile("cp.async.bulk.wait_group %0;\n" ::"n"(N) : "memory"); }

// ------------------------------------------------------------------------------------------------
// tcgen05: TMEM allocation
// ------------------------------------------------------------------------------------------------
FA_DEVICE void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {   // whole warp, ncols power of 2 >= 32
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_result)),
               "r"(ncols)
               : "memory");
}
FA_DEVICE void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
FA_DEVICE void tmem_dealloc(uint32_t taddr, uint32_t ncols) {       // whole warp (the allocating one)
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols) : "memory");
}
FA_DEVICE void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
FA_DEVICE void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
FA_DEVICE void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
FA_DEVICE void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// tcgen05.commit: the mbarrier gets one arrive when all MMAs previously issued by this thread finish.
FA_DEVICE void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(bar))
               : "memory");
}

// ------------------------------------------------------------------------------------------------
// tcgen05.mma (kind::f16: fp16/bf16 inputs, fp32 accumulate). One thread issues.
// ------------------------------------------------------------------------------------------------
// D[tmem] (+)= A[smem] * B[smem]
FA_DEVICE void umma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]   (A: 128 lanes x K/2 32-bit columns of packed 16-bit values)
FA_DEVICE void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Same, with the descriptors given as their LOW words only: for 128-byte-swizzled tiles with SBO = 1024 the high word
// is the constant 0x40004040 (SBO >> 4 | version 1 | swizzle mode 2), so a kernel that addresses many sub-tiles
// only carries 32-bit values around (low word = address >> 4 | (LBO >> 4) << 16; offsets add to the address field).
constexpr uint32_t kDescHiSw128Sbo1024 = 0x40004040u;
FA_DEVICE uint32_t make_smem_desc_lo(uint32_t saddr, uint32_t lbo_bytes) {
  return ((saddr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
FA_DEVICE void umma_ss_lo(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\t"
      "mov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(kDescHiSw128Sbo1024)
      : "memory");
}
FA_DEVICE void umma_ts_lo(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(kDescHiSw128Sbo1024)
      : "memory");
}

// Instruction descriptor for kind::f16 (bit layout: sm_100 UMMA "InstrDescriptor"):
//  [4,6) D format (1 = f32) | [7,10) A format | [10,13) B format (0 = f16, 1 = bf16)
//  [15] A major (0 = K, 1 = MN) | [16] B major | [17,23) N>>3 | [24,29) M>>4
__host__ __device__ constexpr uint32_t make_idesc_f16(bool bf16, int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | ((bf16 ? 1u : 0u) << 7) | ((bf16 ? 1u : 0u) << 10) | ((a_mn_major ? 1u : 0u) << 15) |
         ((b_mn_major ? 1u : 0u) << 16) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}

// Shared-memory matrix descriptor (sm_100 "SmemDescriptor"), 128-byte swizzle:
//  [0,14) start address >> 4 | [16,30) leading byte offset >> 4 | [32,46) stride byte offset >> 4
//  [46,48) version = 1 | [61,64) layout type (2 = SWIZZLE_128B)
// K-major operand tile  (rows = M or N, 64 16-bit elements = 128 B per row, 8-row groups of 1024 B):
//     SBO = 1024 (next 8-row group), LBO unused.
// MN-major operand tile (rows = K, 64 MN elements = 128 B per row):
//     SBO = 1024 (next 8 K-rows), LBO = byte distance to the next 64-element MN group.
FA_DEVICE uint64_t make_smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

// ------------------------------------------------------------------------------------------------
// tcgen05.ld / tcgen05.st, shape 32x32b: thread (lane l of warp w) touches TMEM lane 32*(w%4)+l and
// N consecutive 32-bit columns.  The address carries (lane << 16) | column.
// ------------------------------------------------------------------------------------------------
FA_DEVICE void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
FA_DEVICE void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
FA_DEVICE void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};\n"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
        "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
        "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
FA_DEVICE void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// ------------------------------------------------------------------------------------------------
// misc
// ------------------------------------------------------------------------------------------------
FA_DEVICE float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Packed fp32 pairs (FFMA2 / FADD2 on sm_100): two results per issue slot.  Same rounding as the scalar forms.
FA_DEVICE void ffma2(float& a0, float& a1, float s, float b) {   // (a0, a1) = (a0, a1) * s + b
  uint64_t in, sc, bb, out;
  asm("mov.b64 %0, {%1, %2};" : "=l"(in) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(sc) : "f"(s), "f"(s));
  asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b), "f"(b));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(out) : "l"(in), "l"(sc), "l"(bb));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(out));
}
FA_DEVICE void fadd2(float& a0, float& a1, float b0, float b1) {   // (a0, a1) += (b0, b1)
  uint64_t x, y, out;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(b0), "f"(b1));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(out) : "l"(x), "l"(y));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(out));
}
template <bool kBf16>
FA_DEVICE uint32_t pack2(float lo, float hi) {   // round-to-nearest-even pack of two floats into 16-bit pair
  uint32_t r;
  if (kBf16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  } else {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  }
  return r;
}
// ------------------------------------------------------------------------------------------------
// Epilogue helper: one warp moves its 32 accumulator rows (TMEM lanes, DPAD fp32 columns) to global memory as 16-bit
// rows of `d` elements.  Each lane owns one row in TMEM, which would make every global store touch 32 different
// lines; the rows are therefore staged through 32 * DPAD * 2 bytes of shared memory (XOR-swizzled 16-byte chunks) and
// written back with 16 / 8 lanes per row, i.e. whole 256 / 128-byte rows per store instruction.
// `taddr`: TMEM address of this warp's lane quarter, column 0 of the accumulator.  `have_acc` = false stores zeros.
// ------------------------------------------------------------------------------------------------
template <int DPAD, bool kBf16>
FA_DEVICE void store_acc_rows_coalesced(uint32_t taddr, bool have_acc, float scale, uint32_t stage, uint16_t* grow0,
                                        int64_t row_stride, int valid_rows, int d, int lane) {
  constexpr int kRowBytes = DPAD * 2;
  constexpr int kChunksPerRow = DPAD / 8;             // 16-byte chunks
  constexpr int kRowsPerInstr = 32 / kChunksPerRow;   // 2 (DPAD 128) or 4 (DPAD 64)
#pragma unroll
  for (int c4 = 0; c4 < DPAD / 32; ++c4) {
    uint32_t o[32];
    if (have_acc) {
      tmem_ld32(taddr + c4 * 32, o);
      tmem_wait_ld();
    } else {
#pragma unroll
      for (int c = 0; c < 32; ++c) o[c] = 0u;
    }
#pragma unroll
    for (int v8 = 0; v8 < 4; ++v8) {
      const uint32_t chunk = static_cast<uint32_t>((c4 * 4 + v8) ^ (lane & 7));
      const uint32_t a = pack2<kBf16>(__uint_as_float(o[v8 * 8 + 0]) * scale, __uint_as_float(o[v8 * 8 + 1]) * scale);
      const uint32_t b = pack2<kBf16>(__uint_as_float(o[v8 * 8 + 2]) * scale, __uint_as_float(o[v8 * 8 + 3]) * scale);
      const uint32_t c = pack2<kBf16>(__uint_as_float(o[v8 * 8 + 4]) * scale, __uint_as_float(o[v8 * 8 + 5]) * scale);
      const uint32_t e = pack2<kBf16>(__uint_as_float(o[v8 * 8 + 6]) * scale, __uint_as_float(o[v8 * 8 + 7]) * scale);
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(stage + lane * kRowBytes + chunk * 16), "r"(a), "r"(b),
                   "r"(c), "r"(e)
                   : "memory");
    }
  }
  __syncwarp();
  const int j = lane % kChunksPerRow;
#pragma unroll
  for (int i = 0; i < 32 / kRowsPerInstr; ++i) {
    const int row = i * kRowsPerInstr + lane / kChunksPerRow;
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(stage + row * kRowBytes + ((j ^ (row & 7)) * 16))
                 : "memory");
    if (row < valid_rows && j * 8 < d) *reinterpret_cast<uint4*>(grow0 + row * row_stride + j * 8) = v;
  }
}

FA_DEVICE void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}
FA_DEVICE void named_bar_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}
template <int N>
FA_DEVICE void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(N)); }
template <int N>
FA_DEVICE void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(N)); }

}  // namespace fa
