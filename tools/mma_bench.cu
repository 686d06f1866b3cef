// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, M=128) for the operand forms the attention kernels use.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../hazyresearch_flash-attention_b200/csrc mma_bench.cu -o mma_bench
// Prints cycles per 128 x N x 16 MMA instruction for each variant, with and without concurrent shared-memory stores.
#include <cstdio>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"

using namespace fa;

constexpr int kSub = 16384;

// variant: 0 SS K/K   1 SS K/MN   2 SS MN/MN   3 TS ./MN   4 TS ./K   5 SS K/K N=256   6 bwd mix (S,dP,dV,dK,dQ)  7 SS K/K N=64
__global__ void __launch_bounds__(256, 1) mma_bench_kernel(int variant, int reps, int traffic_warps, int random_data, long long* out, int kind, float* scratch) {
  // random_data: bit0 random operands, bit1 TMEM-load traffic, bit2 + TMEM stores.  kind (when bit1 clear): 0 st.shared, 3 bulk loads
  // global->smem, 4 bulk fp32 reductions smem->global, 5 MUFU.EX2, 6 ld.shared, 7 deep tcgen05.ld (4 x 32 columns in flight)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  __shared__ int stop_flag;
  __shared__ uint64_t tbar[8];
  const int warp = threadIdx.x >> 5;
  // zero operands (4 tiles of 32 KB)
  for (int i = threadIdx.x; i < 4 * 32768 / 4; i += blockDim.x) {
    // random bf16 pairs in (-2, 2): sign + exponent 0x3f.. + random mantissa
    uint32_t h = (i * 2654435761u) ^ (blockIdx.x * 40503u);
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    const uint32_t lo = 0x3f00u | (h & 0x80ffu), hi = 0x3f00u | ((h >> 16) & 0x80ffu);
    reinterpret_cast<uint32_t*>(smem)[i] = (random_data & 1) ? (lo | (hi << 16)) : 0u;
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 8; ++i) mbar_init(&tbar[i], 1);
    fence_mbar_init();
    stop_flag = 0;
  }
  if (warp == 0) {
    tmem_alloc(&tmem_ptr, 512);
    tmem_relinquish();
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  const uint32_t a0 = smem_u32(smem), b0 = a0 + 32768, c0 = a0 + 65536, d0 = a0 + 98304;
  if (warp == 0) {
    if (elect_one()) {
      const uint32_t id_kk = make_idesc_f16(true, 128, 128, false, false);
      const uint32_t id_kmn = make_idesc_f16(true, 128, 128, false, true);
      const uint32_t id_mnmn = make_idesc_f16(true, 128, 128, true, true);
      const uint32_t id_kk256 = make_idesc_f16(true, 128, 256, false, false);
      const uint32_t id_kk64 = make_idesc_f16(true, 128, 64, false, false);
      auto ss_kk = [&](uint32_t d, uint32_t a, uint32_t b, uint32_t id) {
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * kSub + (kk & 3) * 32;
          umma_ss(d, make_smem_desc_sw128(a + off, 16, 1024), make_smem_desc_sw128(b + off, 16, 1024), id, 1);
        }
      };
      auto ss_kmn = [&](uint32_t d, uint32_t a, uint32_t b) {
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * kSub + (kk & 3) * 32;
          umma_ss(d, make_smem_desc_sw128(a + off, 16, 1024), make_smem_desc_sw128(b + kk * 2048, kSub, 1024), id_kmn, 1);
        }
      };
      auto ss_mnmn = [&](uint32_t d, uint32_t a, uint32_t b) {
        for (int kk = 0; kk < 8; ++kk)
          umma_ss(d, make_smem_desc_sw128(a + kk * 2048, kSub, 1024), make_smem_desc_sw128(b + kk * 2048, kSub, 1024), id_mnmn, 1);
      };
      auto ts_mn = [&](uint32_t d, uint32_t a, uint32_t b) {
        for (int kk = 0; kk < 8; ++kk) umma_ts(d, a + kk * 8, make_smem_desc_sw128(b + kk * 2048, kSub, 1024), id_kmn, 1);
      };
      auto ts_k = [&](uint32_t d, uint32_t a, uint32_t b) {
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * kSub + (kk & 3) * 32;
          umma_ts(d, a + kk * 8, make_smem_desc_sw128(b + off, 16, 1024), id_kk, 1);
        }
      };
      const long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        switch (variant) {
          case 0: ss_kk(tm, a0, b0, id_kk); break;
          case 1: ss_kmn(tm, a0, b0); break;
          case 2: ss_mnmn(tm, a0, b0); break;
          case 3: ts_mn(tm, tm + 256, b0); break;
          case 4: ts_k(tm, tm + 256, b0); break;
          case 5: ss_kk(tm, a0, b0, id_kk256); break;   // B tile = 2 x 128 rows: b0 and c0 halves are contiguous enough for timing
          case 6:
            ss_kk(tm, a0, b0, id_kk);          // S^T
            ss_kk(tm + 128, c0, d0, id_kk);    // dP^T
            ts_mn(tm + 256, tm, d0);           // dV
            ss_kmn(tm + 384, a0, b0);          // dK
            ss_mnmn(tm + 128, a0, c0);         // dQ
            break;
          case 7: ss_kk(tm, a0, b0, id_kk64); break;
        }
      }
      umma_commit(&bar);
      mbar_wait(&bar, 0);
      const long long t1 = clock64();
      if (blockIdx.x == 0) out[0] = t1 - t0;
      stop_flag = 1;
      __threadfence_block();
    }
  } else if (warp <= traffic_warps) {
    // concurrent shared-memory store traffic (16 B per lane per store, conflict free) in a scratch region after the operands
    const uint32_t dst = a0 + 131072 + (threadIdx.x & 1023) * 16;
    volatile int* sf = &stop_flag;
    if (random_data & 2) {
      // TMEM traffic instead: every warp reads (bit 2: also writes back) 32 columns of its lane quarter in a loop
      const uint32_t t = tm + (static_cast<uint32_t>((warp & 3) * 32) << 16) + 448;
      uint32_t r[32];
      while (*sf == 0) {
        tmem_ld32(t, r);
        tmem_wait_ld();
        if (random_data & 4) { tmem_st32(t + 32, r); tmem_wait_st(); }
      }
    } else if (kind == 3) {
      // bulk loads global -> shared (8 KB each, private 8 KB slot per warp), like the Q / dO tile loads
      uint8_t* slot = smem + 131072 + (warp - 1) * 2048;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(scratch) + (size_t)blockIdx.x * 65536 + (warp - 1) * 8192;
      uint32_t ph = 0;
      while (*sf == 0) {
        if ((threadIdx.x & 31) == 0) {
          mbar_arrive_expect_tx(&tbar[warp], 2048);
          bulk_load_1d(slot, src, 2048, &tbar[warp]);
          mbar_wait(&tbar[warp], ph);
        }
        ph ^= 1;
        __syncwarp();
      }
    } else if (kind == 4) {
      // bulk fp32 reductions shared -> global (2 KB each), like the dQ drain
      uint8_t* slot = smem + 131072 + (warp - 1) * 2048;
      float* dst = scratch + (size_t)blockIdx.x * 16384 + (warp - 1) * 2048;
      int n = 0;
      while (*sf == 0) {
        if ((threadIdx.x & 31) == 0) {
          bulk_reduce_add_f32(dst + (n & 3) * 512, slot, 2048);
          bulk_commit_group();
          bulk_wait_group_read<2>();
        }
        ++n;
        __syncwarp();
      }
      if ((threadIdx.x & 31) == 0) bulk_wait_group<0>();
    } else if (kind == 5) {
      float x = threadIdx.x * 1e-3f, acc = 0.f;
      while (*sf == 0) {
#pragma unroll
        for (int i = 0; i < 32; ++i) acc += ex2(x + i * 1e-4f);
        x += 1e-5f;
      }
      if (acc == 1234.5f) out[1] = 1;
    } else if (kind == 6) {
      const uint32_t srcs = a0 + 131072 + (threadIdx.x & 1023) * 16;
      uint32_t acc = 0;
      while (*sf == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint32_t x, y, z, w;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(srcs));
          acc += x + y + z + w;
        }
      }
      if (acc == 0x12345u) out[1] = 1;
    } else if (kind == 7) {
      const uint32_t t = tm + (static_cast<uint32_t>((warp & 3) * 32) << 16) + 256;
      uint32_t r0[32], r1[32], r2[32], r3[32];
      uint32_t acc = 0;
      while (*sf == 0) {
        tmem_ld32(t, r0); tmem_ld32(t + 32, r1); tmem_ld32(t + 64, r2); tmem_ld32(t + 96, r3);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc += r0[i] ^ r1[i] ^ r2[i] ^ r3[i];
      }
      if (acc == 0x12345u) out[1] = 1;
    } else
    while (*sf == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};\n" ::"r"(dst), "r"(i) : "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tm, 512);
  }
}

int main() {
  long long* out;
  cudaMalloc(&out, 16);
  const int smem = 4 * 32768 + 16384 + 1024;
  cudaFuncSetAttribute(mma_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"SS K/K N128", "SS K/MN N128", "SS MN/MN N128", "TS ./MN N128", "TS ./K N128", "SS K/K N256", "bwd mix (5 tiles)", "SS K/K N64"};
  const int reps = 200;
  float* scratch;
  cudaMalloc(&scratch, (size_t)148 * 65536 * 4);
  cudaMemset(scratch, 0, (size_t)148 * 65536 * 4);
  const char* kinds[] = {"st.shared", "tmem ld", "tmem ld+st", "bulk load", "bulk reduce", "MUFU.EX2", "ld.shared", "deep tmem ld"};
  struct Cfg { int warps, random, kind; };
  const Cfg cfgs[] = {{0, 1, 0}, {7, 1, 0}, {7, 3, 1}, {7, 7, 2}, {4, 1, 3}, {7, 1, 3}, {4, 1, 4}, {7, 1, 4}, {7, 1, 5}, {7, 1, 6}, {4, 1, 7}, {7, 1, 7}};
  for (const Cfg& c : cfgs) {
    for (int v : {0, 2, 3, 6}) {
      mma_bench_kernel<<<148, 256, smem>>>(v, reps, c.warps, c.random, out, c.kind, scratch);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("variant %d failed: %s\n", v, cudaGetErrorString(e)); return 1; }
      long long cyc;
      cudaMemcpy(&cyc, out, 8, cudaMemcpyDeviceToHost);
      const int n_instr = reps * 8 * (v == 6 ? 5 : 1);
      printf("traffic: %d warps of %-12s  %-18s  %8.1f clk / MMA instr\n", c.warps, c.warps ? kinds[c.kind] : "-", names[v], (double)cyc / n_instr);
    }
  }
  return 0;
}
