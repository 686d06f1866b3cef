// Micro-benchmark: chip-wide throughput of fp32 reductions into global memory (the dQ accumulation path of the backward).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../hazyresearch_flash-attention_b200/csrc red_bench.cu -o red_bench
// mode 0: cp.reduce.async.bulk.add.f32 (smem -> global)   mode 1: cp.async.bulk plain store (smem -> global)
// mode 2: red.global.add.v4.f32 from registers (each lane 16 B, lanes contiguous)   mode 3: st.global.v4 (same pattern)
#include <cstdio>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"

using namespace fa;

__global__ void __launch_bounds__(128, 1) red_bench_kernel(float* dst, int mode, int op_bytes, int ops_per_warp,
                                                           long long window_bytes, int inflight, int share, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 1.0f;
  fence_proxy_async();
  __syncthreads();
  // every warp streams over its own slice of the CTA's window
  const long long cta_base = (long long)(blockIdx.x / share) * window_bytes;   // `share` CTAs hit the same addresses
  const long long warp_span = window_bytes / 4;
  uint8_t* gbase = reinterpret_cast<uint8_t*>(dst) + cta_base + warp * warp_span;
  const long long t0 = clock64();
  if (mode <= 1) {
    if (lane == 0) {
      long long off = 0;
      for (int i = 0; i < ops_per_warp; ++i) {
        void* src = smem + warp * 16384;
        if (mode == 0) bulk_reduce_add_f32(gbase + off, src, op_bytes);
        else asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gbase + off), "r"(smem_u32(src)), "r"(op_bytes) : "memory");
        bulk_commit_group();
        if (inflight == 1) bulk_wait_group_read<0>();
        else if (inflight == 2) bulk_wait_group_read<1>();
        else if (inflight == 4) bulk_wait_group_read<3>();
        else bulk_wait_group_read<7>();
        off += op_bytes;
        if (off + op_bytes > warp_span) off = 0;
      }
      bulk_wait_group<0>();
    }
  } else {
    long long off = 0;
    const int per_instr = 512;   // 32 lanes x 16 B
    const int n = ops_per_warp * (op_bytes / per_instr);
    for (int i = 0; i < n; ++i) {
      float* g = reinterpret_cast<float*>(gbase + off) + lane * 4;
      if (mode == 2) asm volatile("red.global.add.v4.f32 [%0], {%1, %1, %1, %1};\n" ::"l"(g), "f"(1.0f) : "memory");
      else asm volatile("st.global.v4.f32 [%0], {%1, %1, %1, %1};\n" ::"l"(g), "f"(1.0f) : "memory");
      off += per_instr;
      if (off + per_instr > warp_span) off = 0;
    }
    __threadfence();
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

int main() {
  const long long max_window = 4ll << 20;   // per CTA
  float* dst;
  cudaMalloc(&dst, 148 * max_window);
  cudaMemset(dst, 0, 148 * max_window);
  long long* out;
  cudaMalloc(&out, 148 * 8);
  const int smem = 65536 + 1024;
  cudaFuncSetAttribute(red_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"bulk reduce add.f32", "bulk store", "red.global.v4.f32", "st.global.v4"};
  struct Cfg { int mode, op_bytes, inflight; long long window; int share; };
  const Cfg cfgs[] = {
      {0, 2048, 2, 1 << 20}, {0, 2048, 8, 1 << 20}, {0, 8192, 2, 1 << 20}, {0, 8192, 8, 1 << 20}, {0, 16384, 8, 1 << 20},
      {0, 2048, 8, 64 << 10}, {0, 8192, 8, 4 << 20},
      {1, 2048, 8, 1 << 20}, {1, 8192, 8, 1 << 20},
      {2, 2048, 0, 1 << 20}, {2, 2048, 0, 64 << 10}, {3, 2048, 0, 1 << 20},
      {0, 2048, 8, 64 << 10, 2}, {0, 2048, 8, 64 << 10, 8}, {0, 2048, 8, 64 << 10, 32}, {0, 2048, 8, 2 << 20, 32},
      {0, 2048, 2, 64 << 10, 32}, {1, 2048, 8, 64 << 10, 32},
  };
  for (const Cfg& c : cfgs) {
    const int ops = (8 << 20) / c.op_bytes / 4;   // 8 MB per CTA in total
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    red_bench_kernel<<<148, 128, smem>>>(dst, c.mode, c.op_bytes, ops, c.window, c.inflight, c.share ? c.share : 1, out);   // warm-up
    cudaEventRecord(e0);
    red_bench_kernel<<<148, 128, smem>>>(dst, c.mode, c.op_bytes, ops, c.window, c.inflight, c.share ? c.share : 1, out);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(e)); return 1; }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long cyc[148];
    cudaMemcpy(cyc, out, sizeof(cyc), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < 148; ++i) mx = cyc[i] > mx ? cyc[i] : mx;
    const double bytes = 148.0 * 4 * ops * c.op_bytes;
    printf("%-22s op %5d B  inflight %d  share %2d  window %4lld KB : %7.1f GB/s  %6.1f B/clk/SM  (%.3f ms)\n", names[c.mode], c.op_bytes,
           c.inflight, c.share ? c.share : 1, c.window >> 10, bytes / ms / 1e6, (double)4 * ops * c.op_bytes / mx, ms);
  }
  return 0;
}
