// Micro-benchmark: tcgen05.ld / tcgen05.st bandwidth (TMEM <-> registers) with 4, 8 and 12 warps.
#include <cstdio>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"
using namespace fa;

__global__ void __launch_bounds__(512, 1) tmem_bench_kernel(int mode, int nwarps, int reps, long long* out, uint32_t* sink) {
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc(&tmem_ptr, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t t = tmem_ptr + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (warp >> 2) * 128;
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    for (int r = 0; r < reps; ++r) {
      if (mode == 0) {          // 4 x (ld x32) then one wait : 128 columns = 16 KB per warp
        uint32_t a[32], b[32], c[32], d[32];
        tmem_ld32(t, a); tmem_ld32(t + 32, b); tmem_ld32(t + 64, c); tmem_ld32(t + 96, d);
        tmem_wait_ld();
        acc += a[0] ^ b[1] ^ c[2] ^ d[3];
      } else if (mode == 1) {   // ld x32 + wait each
        for (int j = 0; j < 4; ++j) { uint32_t a[32]; tmem_ld32(t + j * 32, a); tmem_wait_ld(); acc += a[j]; }
      } else if (mode == 2) {   // st x32 x4 + wait
        uint32_t a[32];
        for (int j = 0; j < 32; ++j) a[j] = r + j;
        tmem_st32(t, a); tmem_st32(t + 32, a); tmem_st32(t + 64, a); tmem_st32(t + 96, a);
        tmem_wait_st();
      } else {                  // ld x16 x8
        for (int j = 0; j < 8; ++j) { uint32_t a[16]; tmem_ld16(t + j * 16, a); tmem_wait_ld(); acc += a[j]; }
      }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  if (acc == 0x12345678) sink[0] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_ptr, 512); }
}

int main() {
  long long* out; uint32_t* sink;
  cudaMalloc(&out, 8); cudaMalloc(&sink, 4);
  const char* names[] = {"ld x32 x4, one wait", "ld x32 + wait each", "st x32 x4, one wait", "ld x16 + wait each"};
  const int reps = 2000;
  for (int mode = 0; mode < 4; ++mode)
    for (int nw : {1, 4, 8, 16}) {
      tmem_bench_kernel<<<148, 512>>>(mode, nw, reps, out, sink);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("failed\n"); return 1; }
      long long cyc; cudaMemcpy(&cyc, out, 8, cudaMemcpyDeviceToHost);
      const double bytes = (double)nw * reps * 128 * 32 * 4;
      printf("%-22s warps %2d : %7.1f B/clk/SM   (%.0f clk per 64 KB tile)\n", names[mode], nw, bytes / cyc, 65536.0 / (bytes / cyc));
    }
  return 0;
}
