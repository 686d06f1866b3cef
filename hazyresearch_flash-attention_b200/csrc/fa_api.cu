// C ABI of libfa_b200.so (see include/fa_b200.h).  Host-side launcher: argument validation, device
// check, workspace carving, kernel dispatch.  B200 equivalent of mha_fwd / set_params_fprop / run_fmha_fprop
// (csrc/flash_attn/fmha_api.cpp:38-242, src/fmha_fprop_kernel_dispatch.cu:90-134) -- but torch-free, never
// allocating, never synchronising, and reporting errors by return code instead of exit(1)
// (src/fmha_utils.h:37-48).
#include <cuda.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>
#include <vector>

#include "fa_common.cuh"
#include "fa_kernels.h"

namespace fa {

namespace {

thread_local char g_last_error[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char* what) {
  return fail(FA_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline uint64_t round_up(uint64_t x, uint64_t m) { return (x + m - 1) / m * m; }

int check_device() {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  static std::atomic<int> cached_major[64];   // 0 = not queried yet; concurrent first calls just query twice
  int major = dev < 64 ? cached_major[dev].load(std::memory_order_relaxed) : 0;
  if (major == 0) {
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    if (dev < 64) cached_major[dev].store(major, std::memory_order_relaxed);
  }
  if (major != 10)
    return fail(FA_ERR_UNSUPPORTED_DEVICE, "libfa_b200 needs a compute capability 10.x (sm_100a) device, found %d.x", major);
  return FA_OK;
}

// Shape / dtype / stride checks corresponding to the TORCH_CHECKs of fmha_api.cpp:127-170.
int validate_common(const fa_params* p, bool bwd, bool block) {
  if (p == nullptr) return fail(FA_ERR_INVALID_ARGUMENT, "params is NULL");
  if (p->b <= 0) return fail(FA_ERR_INVALID_ARGUMENT, "batch size must be positive (got %d)", p->b);
  if (p->h <= 0) return fail(FA_ERR_INVALID_ARGUMENT, "number of heads must be positive (got %d)", p->h);
  if (p->d <= 0 || p->d % 8 != 0 || p->d > 128)
    return fail(FA_ERR_INVALID_ARGUMENT, "head dim must be a multiple of 8 and <= 128 (got %d)", p->d);
  if (p->total_q < 0 || p->total_k < 0) return fail(FA_ERR_INVALID_ARGUMENT, "negative token count");
  if (p->max_seqlen_q < 0 || p->max_seqlen_k < 0) return fail(FA_ERR_INVALID_ARGUMENT, "negative max_seqlen");
  if (!(p->p_dropout >= 0.f && p->p_dropout < 1.f))
    return fail(FA_ERR_INVALID_ARGUMENT, "p_dropout must be in [0, 1) (got %f)", (double)p->p_dropout);
  if (p->is_bf16 != 0 && p->is_bf16 != 1) return fail(FA_ERR_INVALID_ARGUMENT, "is_bf16 must be 0 or 1");
  if (p->impl < FA_IMPL_AUTO || p->impl > FA_IMPL_SIMT) return fail(FA_ERR_INVALID_ARGUMENT, "unknown impl %d", p->impl);
  if (p->cu_seqlens_q == nullptr || p->cu_seqlens_k == nullptr)
    return fail(FA_ERR_INVALID_ARGUMENT, "cu_seqlens_q / cu_seqlens_k must be non-NULL int32 device arrays");
  if (p->seqlen_q_rounded < p->max_seqlen_q)
    return fail(FA_ERR_INVALID_ARGUMENT, "seqlen_q_rounded (%d) < max_seqlen_q (%d)", p->seqlen_q_rounded, p->max_seqlen_q);
  if (p->softmax_lse == nullptr) return fail(FA_ERR_INVALID_ARGUMENT, "softmax_lse is NULL");
  if (p->s_dmask != nullptr && p->seqlen_k_rounded < p->max_seqlen_k)
    return fail(FA_ERR_INVALID_ARGUMENT, "seqlen_k_rounded (%d) < max_seqlen_k (%d)", p->seqlen_k_rounded, p->max_seqlen_k);

  struct T { const void* ptr; int64_t rs, hs; const char* name; bool need; };
  const T ts[] = {
      {p->q, p->q_row_stride, p->q_head_stride, "q", true},
      {p->k, p->k_row_stride, p->k_head_stride, "k", true},
      {p->v, p->v_row_stride, p->v_head_stride, "v", true},
      {p->o, p->o_row_stride, p->o_head_stride, "o", true},
      {p->dout, p->do_row_stride, p->do_head_stride, "dout", bwd},
      {p->dq, p->dq_row_stride, p->dq_head_stride, "dq", bwd && !(p->flags & FA_FLAG_KEEP_DQ_ACCUM)},
      {p->dk, p->dk_row_stride, p->dk_head_stride, "dk", bwd},
      {p->dv, p->dv_row_stride, p->dv_head_stride, "dv", bwd},
  };
  for (const T& t : ts) {
    if (!t.need) continue;
    const bool is_k_side = (t.name[0] == 'k' || t.name[0] == 'v' || (t.name[0] == 'd' && (t.name[1] == 'k' || t.name[1] == 'v')));
    const int total = is_k_side ? p->total_k : p->total_q;
    if (t.ptr == nullptr && total > 0) return fail(FA_ERR_INVALID_ARGUMENT, "%s is NULL", t.name);
    if (!aligned16(t.ptr)) return fail(FA_ERR_INVALID_ARGUMENT, "%s must be 16-byte aligned", t.name);
    if (t.rs % 8 != 0 || t.hs % 8 != 0 || t.rs < p->d || t.hs < p->d)
      return fail(FA_ERR_INVALID_ARGUMENT, "%s strides (row %lld, head %lld) must be multiples of 8 elements and >= d",
                  t.name, (long long)t.rs, (long long)t.hs);
  }
  if (bwd && p->softmax_d == nullptr) return fail(FA_ERR_INVALID_ARGUMENT, "softmax_d is NULL");
  if (p->flags & ~(FA_FLAG_WS_ZEROED | FA_FLAG_KEEP_DQ_ACCUM | FA_FLAG_DETERMINISTIC)) return fail(FA_ERR_INVALID_ARGUMENT, "unknown flags 0x%x", p->flags);
  if ((p->flags & FA_FLAG_DETERMINISTIC) && (!bwd || block || p->impl == FA_IMPL_SIMT || (p->flags & FA_FLAG_KEEP_DQ_ACCUM)))
    return fail(FA_ERR_INVALID_ARGUMENT, "FA_FLAG_DETERMINISTIC applies to fa_bwd with the tcgen05 kernels, without a block mask and without "
                                         "FA_FLAG_KEEP_DQ_ACCUM");
  if ((p->flags & FA_FLAG_KEEP_DQ_ACCUM) && (!bwd || block || p->impl == FA_IMPL_SIMT))
    return fail(FA_ERR_INVALID_ARGUMENT, "FA_FLAG_KEEP_DQ_ACCUM applies to fa_bwd with the tcgen05 kernels only");
  if ((p->flags & FA_FLAG_KEEP_DQ_ACCUM) && !(p->flags & FA_FLAG_WS_ZEROED))
    return fail(FA_ERR_INVALID_ARGUMENT, "FA_FLAG_KEEP_DQ_ACCUM needs FA_FLAG_WS_ZEROED: the caller owns the workspace and clears it once "
                                         "(without the flag the library would clear the running sum on every call)");
  if (p->alibi_slopes != nullptr && block)
    return fail(FA_ERR_INVALID_ARGUMENT, "blocksparse attention does not take an attention bias");
  if (p->impl == FA_IMPL_SIMT && (p->b > 65535 || p->h > 65535))
    return fail(FA_ERR_INVALID_ARGUMENT, "the SIMT cross-check kernels support at most 65535 sequences / heads");
  if (block) {
    if (p->blockmask == nullptr) return fail(FA_ERR_INVALID_ARGUMENT, "blockmask is NULL");
    if (p->blockmask_nrow <= 0 || p->blockmask_ncol <= 0)
      return fail(FA_ERR_INVALID_ARGUMENT, "blockmask shape must be positive (got %d x %d)", p->blockmask_nrow, p->blockmask_ncol);
    if (p->is_causal) return fail(FA_ERR_INVALID_ARGUMENT, "blocksparse attention does not support causal (flash_blocksparse_attn_interface.py:21)");
  } else if (p->blockmask != nullptr) {
    return fail(FA_ERR_INVALID_ARGUMENT, "blockmask given to a dense entry point");
  }
  return FA_OK;
}

__global__ void expand_blockmask_kernel(const int32_t* __restrict__ conv, int nrow, int ncol, uint8_t* __restrict__ dense) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // over (ncol, nrow)
  if (idx >= nrow * ncol) return;
  const int c = idx / nrow;
  const int val = conv[idx];
  if (val >= 0) {
    const int r = val >> 2;
    if (r < nrow) dense[r * ncol + c] = 1;
  }
}

template <typename V>
__global__ void gather_rows_kernel(V* __restrict__ dst, const V* __restrict__ src, const int64_t* __restrict__ idx,
                                   int64_t n, int64_t row_vecs) {
  // one CTA row-group per blockIdx.y-strided index; threads stride over the row's vectors
  for (int64_t i = blockIdx.x; i < n; i += gridDim.x) {
    const int64_t r = idx[i];
    V* d = dst + i * row_vecs;
    if (r < 0) {                                   // padding entry of a sync-free index list (fa_varlen_index): a zero row
      for (int64_t c = threadIdx.x; c < row_vecs; c += blockDim.x) d[c] = V{};
      continue;
    }
    const V* s = src + r * row_vecs;
    for (int64_t c = threadIdx.x; c < row_vecs; c += blockDim.x) d[c] = __ldg(s + c);
  }
}
template <typename V>
__global__ void scatter_rows_kernel(V* __restrict__ dst, const V* __restrict__ src, const int64_t* __restrict__ idx,
                                    int64_t n, int64_t row_vecs) {
  for (int64_t i = blockIdx.x; i < n; i += gridDim.x) {
    const int64_t r = idx[i];
    if (r < 0) continue;                           // padding entry: no destination row
    const V* s = src + i * row_vecs;
    V* d = dst + r * row_vecs;
    for (int64_t c = threadIdx.x; c < row_vecs; c += blockDim.x) d[c] = __ldg(s + c);
  }
}

// dst[idx[i], :] += src[i, :] (indices unique, negative = skip): the backward of IndexFirstAxisResidual
// (flash_attn/bert_padding.py:67-96, `grad_residual.index_add`).  kind 0 = fp16, 1 = bf16, 2 = fp32; 16 bytes per thread.
template <int kKind>
__global__ void scatter_add_rows_kernel(void* __restrict__ dst_, const void* __restrict__ src_, const int64_t* __restrict__ idx,
                                        int64_t n, int64_t row_vecs) {
  uint4* dst = reinterpret_cast<uint4*>(dst_);
  const uint4* src = reinterpret_cast<const uint4*>(src_);
  for (int64_t i = blockIdx.x; i < n; i += gridDim.x) {
    const int64_t r = idx[i];
    if (r < 0) continue;
    for (int64_t c = threadIdx.x; c < row_vecs; c += blockDim.x) {
      uint4 a = dst[r * row_vecs + c];
      const uint4 b = __ldg(src + i * row_vecs + c);
      if (kKind == 2) {
        float* x = reinterpret_cast<float*>(&a);
        const float* y = reinterpret_cast<const float*>(&b);
#pragma unroll
        for (int t = 0; t < 4; ++t) x[t] += y[t];
      } else {
        using T = typename Elem<kKind == 1>::T;
        T* x = reinterpret_cast<T*>(&a);
        const T* y = reinterpret_cast<const T*>(&b);
#pragma unroll
        for (int t = 0; t < 8; ++t) x[t] = Elem<kKind == 1>::from_float(Elem<kKind == 1>::to_float(x[t]) + Elem<kKind == 1>::to_float(y[t]));
      }
      dst[r * row_vecs + c] = a;
    }
  }
}

// ---- fa_varlen_index: (batch, seqlen) 0/1 byte mask -> indices of the valid tokens, cu_seqlens, max_seqlen, all on the
// device (the reference does nonzero() + .item(): a host round trip per call, flash_attn/bert_padding.py:99-119). ----
__global__ void __launch_bounds__(256) varlen_count_kernel(const uint8_t* __restrict__ mask, int s, int32_t* __restrict__ cu_seqlens) {
  __shared__ int warp_sums[8];
  const uint8_t* row = mask + (int64_t)blockIdx.x * s;
  int n = 0;
  for (int j = threadIdx.x; j < s; j += 256) n += row[j] != 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += warp_sums[w];
    cu_seqlens[blockIdx.x + 1] = t;                // the length for now; varlen_scan_kernel turns it into the prefix sum
  }
}
__global__ void __launch_bounds__(1024) varlen_scan_kernel(int32_t* __restrict__ cu_seqlens, int b, int32_t* __restrict__ stats) {
  __shared__ int warp_tot[32];
  __shared__ int warp_max[32];
  __shared__ int carry_s, max_s;
  if (threadIdx.x == 0) { carry_s = 0; max_s = 0; cu_seqlens[0] = 0; }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < b; base += 1024) {
    const int i = base + threadIdx.x;
    const int len = i < b ? cu_seqlens[i + 1] : 0;
    int incl = len, mx = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
      mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (lane == 31) warp_tot[warp] = incl;
    if (lane == 0) warp_max[warp] = mx;
    __syncthreads();
    int before = carry_s;
    for (int w = 0; w < warp; ++w) before += warp_tot[w];
    if (i < b) cu_seqlens[i + 1] = before + incl;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0, m = max_s;
      for (int w = 0; w < 32; ++w) { t += warp_tot[w]; m = max(m, warp_max[w]); }
      carry_s += t;
      max_s = m;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { stats[0] = max_s; stats[1] = carry_s; }
}
__global__ void __launch_bounds__(256) varlen_indices_kernel(const uint8_t* __restrict__ mask, int s, const int32_t* __restrict__ cu_seqlens,
                                                             int64_t* __restrict__ indices) {
  __shared__ int warp_cnt[8];
  __shared__ int carry_s;
  const int64_t row0 = (int64_t)blockIdx.x * s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = cu_seqlens[blockIdx.x];
  __syncthreads();
  for (int base = 0; base < s; base += 256) {
    const int j = base + threadIdx.x;
    const bool on = j < s && mask[row0 + j] != 0;
    const uint32_t ball = __ballot_sync(0xffffffffu, on);
    if (lane == 0) warp_cnt[warp] = __popc(ball);
    __syncthreads();
    int before = carry_s;
    for (int w = 0; w < warp; ++w) before += warp_cnt[w];
    if (on) indices[before + __popc(ball & ((1u << lane) - 1u))] = row0 + j;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < 8; ++w) t += warp_cnt[w];
      carry_s += t;
    }
    __syncthreads();
  }
}

// Ring / split-KV merge of two partial attention results over the same queries (the reference's inter-K-block rule,
// src/fmha_fprop_kernel_1xN.h:606-650, lifted to whole blocks):
//     lse' = log(exp(lse) + exp(lse_blk));   o' = o * exp(lse - lse') + o_blk * exp(lse_blk - lse')
// o_acc fp32 and o_blk 16 bit are (rows, h, d) contiguous with row = bi * batch_rows + t, t < seqlen (batch_rows >= seqlen
// lets a call address one half of every sequence); lse_acc (b, h, acc_stride), lse_blk (b, h, blk_stride).  In place:
// one thread per 8 head-dim elements, the d/8 threads of a (row, head) pair are padded to a power of two so that they
// always sit in ONE warp -- all of them read lse_acc, the warp synchronises, then lane c8 = 0 overwrites it.
template <bool kBf16>
__global__ void __launch_bounds__(256) merge_partials_kernel(float* __restrict__ o_acc, float* __restrict__ lse_acc,
                                                             const void* __restrict__ o_blk_, const float* __restrict__ lse_blk,
                                                             int b, int h, int d, int seqlen, int64_t batch_rows, int acc_stride,
                                                             int blk_stride, int d8p, int first) {
  using T = typename Elem<kBf16>::T;
  const int d8 = d >> 3;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)b * seqlen * h * d8p;
  const bool in_range = idx < total;
  const int c8 = static_cast<int>(idx % d8p);
  const int hh = static_cast<int>((idx / d8p) % h);
  const int64_t row = idx / ((int64_t)d8p * h);
  const int bi = in_range ? static_cast<int>(row / seqlen) : 0, t = in_range ? static_cast<int>(row % seqlen) : 0;
  const int64_t ia = ((int64_t)bi * h + hh) * acc_stride + t, ib = ((int64_t)bi * h + hh) * blk_stride + t;
  float wa = 0.f, wb = 0.f, ln = -INFINITY;
  if (in_range) {
    const float lb = lse_blk[ib];
    const float la = first ? -INFINITY : lse_acc[ia];
    const float mx = fmaxf(la, lb);
    if (mx != -INFINITY) {
      const float ea = __expf(la - mx), eb = __expf(lb - mx);
      ln = mx + __logf(ea + eb);
      wa = ea / (ea + eb);
      wb = eb / (ea + eb);
    }
  }
  __syncwarp();
  if (!in_range) return;
  if (c8 == 0) lse_acc[ia] = ln;
  if (c8 >= d8) return;
  const int64_t off = (((int64_t)bi * batch_rows + t) * h + hh) * d + c8 * 8;
  const uint4 raw = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(o_blk_) + off));
  const T* e = reinterpret_cast<const T*>(&raw);
  float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
  if (!first) {
    a0 = *reinterpret_cast<const float4*>(o_acc + off);
    a1 = *reinterpret_cast<const float4*>(o_acc + off + 4);
  }
  a0.x = a0.x * wa + Elem<kBf16>::to_float(e[0]) * wb;
  a0.y = a0.y * wa + Elem<kBf16>::to_float(e[1]) * wb;
  a0.z = a0.z * wa + Elem<kBf16>::to_float(e[2]) * wb;
  a0.w = a0.w * wa + Elem<kBf16>::to_float(e[3]) * wb;
  a1.x = a1.x * wa + Elem<kBf16>::to_float(e[4]) * wb;
  a1.y = a1.y * wa + Elem<kBf16>::to_float(e[5]) * wb;
  a1.z = a1.z * wa + Elem<kBf16>::to_float(e[6]) * wb;
  a1.w = a1.w * wa + Elem<kBf16>::to_float(e[7]) * wb;
  *reinterpret_cast<float4*>(o_acc + off) = a0;
  *reinterpret_cast<float4*>(o_acc + off + 4) = a1;
}

// dst (fp32) += src (16 bit) over n_batches runs of run_elems contiguous elements (run_elems % 8 == 0): the ring backward
// adds each step's dQ / dK / dV block into its fp32 accumulators.  HBM-bound, 8 elements per thread, grid-stride.
template <bool kBf16>
__global__ void __launch_bounds__(256) accumulate_f32_kernel(float* __restrict__ dst, const void* __restrict__ src_, int64_t n_batches,
                                                             int64_t run_elems, int64_t dst_batch_stride, int64_t src_batch_stride) {
  using T = typename Elem<kBf16>::T;
  const int64_t run8 = run_elems >> 3, total = n_batches * run8;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t bi = idx / run8, e8 = idx - bi * run8;
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(src_) + bi * src_batch_stride + e8 * 8));
    const T* e = reinterpret_cast<const T*>(&raw);
    float4* d4 = reinterpret_cast<float4*>(dst + bi * dst_batch_stride + e8 * 8);
    float4 a0 = d4[0], a1 = d4[1];
    a0.x += Elem<kBf16>::to_float(e[0]); a0.y += Elem<kBf16>::to_float(e[1]);
    a0.z += Elem<kBf16>::to_float(e[2]); a0.w += Elem<kBf16>::to_float(e[3]);
    a1.x += Elem<kBf16>::to_float(e[4]); a1.y += Elem<kBf16>::to_float(e[5]);
    a1.z += Elem<kBf16>::to_float(e[6]); a1.w += Elem<kBf16>::to_float(e[7]);
    d4[0] = a0;
    d4[1] = a1;
  }
}

// Cross-GPU signalling for the peer-memory ring (flash_attn/ring_attention.py): monotone 64-bit counters in memory that a
// peer process has mapped through CUDA IPC.  The writer's stream has completed everything the value announces (stream
// order), the system-scope fence publishes it; the waiter spins on its own stream until the counter has reached the value.
__global__ void flag_write_kernel(uint64_t* flag, uint64_t value) {
  __threadfence_system();
  *reinterpret_cast<volatile uint64_t*>(flag) = value;
  __threadfence_system();
}
__global__ void flag_wait_geq_kernel(const uint64_t* flag, uint64_t value, uint64_t timeout_ns) {
  const volatile uint64_t* f = reinterpret_cast<const volatile uint64_t*>(flag);
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (*f < value) {
    __nanosleep(200);
    if ((++spins & 1023u) == 0) {      // a lost peer must surface as a launch failure, never as a hung GPU
      uint64_t now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > timeout_ns) { asm volatile("trap;\n"); }
    }
  }
  __threadfence_system();
}

__global__ void philox_advance_kernel(uint64_t* state, uint64_t* out, uint64_t increment) {
  const uint64_t v = *state;
  *out = v;
  *state = v + increment;
}

// Rotary position embedding applied to the first `n_rot` parts (q, k) of a packed (rows, n_parts, h, d) tensor in ONE
// pass (the reference does it in pure PyTorch: flash_attn/rotary.py:21-42 -- two elementwise products, a rotate_half
// copy and an add per tensor -- followed by a torch.stack copy of q, k, v, flash_attn/flash_attention.py:105-110).
// Interleaved pairs: (x0, x1) -> (x0 cos - x1 sin, x1 cos + x0 sin); `sign` = -1 applies the transpose (backward).
// cos / sin: fp32 (seqlen, d/2), position = row % seqlen.  Parts >= n_rot are copied when in != out.  fp32 math, one
// rounding.  HBM-bound: 16-byte vectors, grid-stride over (row, part, head, 8-element group).
template <bool kBf16>
__global__ void __launch_bounds__(256) rotary_qk_kernel(const void* __restrict__ in_, void* __restrict__ out_, const float* __restrict__ cos_t,
                                                        const float* __restrict__ sin_t, int64_t n_rows, int seqlen, int n_parts,
                                                        int n_rot, int h, int d, int64_t in_row_stride, int64_t out_row_stride,
                                                        float sign) {
  using T = typename Elem<kBf16>::T;
  const T* in = reinterpret_cast<const T*>(in_);
  T* out = reinterpret_cast<T*>(out_);
  const int d8 = d >> 3;
  const int64_t per_row = (int64_t)n_parts * h * d8;
  const int64_t total = n_rows * per_row;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / per_row;
    const int r = static_cast<int>(idx - row * per_row);
    const int c8 = r % d8;
    const int part = r / (d8 * h);
    const int64_t inner = (int64_t)(r / d8) * d + c8 * 8;       // (part * h + head) * d + column
    uint4 raw = __ldg(reinterpret_cast<const uint4*>(in + row * in_row_stride + inner));
    if (part < n_rot) {
      const int pos = static_cast<int>(row % seqlen);
      const float4 c = __ldg(reinterpret_cast<const float4*>(cos_t + (int64_t)pos * (d >> 1) + c8 * 4));
      const float4 sn = __ldg(reinterpret_cast<const float4*>(sin_t + (int64_t)pos * (d >> 1) + c8 * 4));
      const float cs[4] = {c.x, c.y, c.z, c.w};
      const float ss[4] = {sn.x * sign, sn.y * sign, sn.z * sign, sn.w * sign};
      T* e = reinterpret_cast<T*>(&raw);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float x0 = Elem<kBf16>::to_float(e[2 * j]), x1 = Elem<kBf16>::to_float(e[2 * j + 1]);
        e[2 * j] = Elem<kBf16>::from_float(x0 * cs[j] - x1 * ss[j]);
        e[2 * j + 1] = Elem<kBf16>::from_float(x1 * cs[j] + x0 * ss[j]);
      }
    }
    *reinterpret_cast<uint4*>(out + row * out_row_stride + inner) = raw;
  }
}

template <bool kGather>
int move_rows(void* dst, const void* src, const int64_t* indices, int64_t n, int64_t row_bytes, void* stream) {
  if (n < 0 || row_bytes < 0) return fail(FA_ERR_INVALID_ARGUMENT, "negative size");
  if (n == 0 || row_bytes == 0) return FA_OK;
  if (dst == nullptr || src == nullptr || indices == nullptr) return fail(FA_ERR_INVALID_ARGUMENT, "NULL pointer");
  if (row_bytes % 4 != 0) return fail(FA_ERR_INVALID_ARGUMENT, "row_bytes must be a multiple of 4 (got %lld)", (long long)row_bytes);
  int rc = check_device();
  if (rc != FA_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool v16 = row_bytes % 16 == 0 && aligned16(dst) && aligned16(src);
  const int64_t vecs = v16 ? row_bytes / 16 : row_bytes / 4;
  int threads = 32;
  while (threads < 256 && threads < vecs) threads *= 2;
  const unsigned grid = static_cast<unsigned>(n < 148 * 16 ? n : 148 * 16);
  if (v16) {
    if (kGather) gather_rows_kernel<uint4><<<grid, threads, 0, st>>>((uint4*)dst, (const uint4*)src, indices, n, vecs);
    else scatter_rows_kernel<uint4><<<grid, threads, 0, st>>>((uint4*)dst, (const uint4*)src, indices, n, vecs);
  } else {
    if (kGather) gather_rows_kernel<uint32_t><<<grid, threads, 0, st>>>((uint32_t*)dst, (const uint32_t*)src, indices, n, vecs);
    else scatter_rows_kernel<uint32_t><<<grid, threads, 0, st>>>((uint32_t*)dst, (const uint32_t*)src, indices, n, vecs);
  }
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, kGather ? "gather_rows launch" : "scatter_rows launch");
  return FA_OK;
}

uint64_t blockmask_bytes(const fa_params* p) {
  return round_up(static_cast<uint64_t>(p->blockmask_nrow) * p->blockmask_ncol, 256);
}

// Builds the dense byte mask in the head of the workspace.
int prepare_blockmask(const fa_params* p, cudaStream_t st, BlockMaskView* bm) {
  const uint64_t need = blockmask_bytes(p);
  if (p->workspace == nullptr || p->workspace_bytes < need)
    return fail(FA_ERR_WORKSPACE, "workspace too small: need %llu bytes, have %llu", (unsigned long long)need,
                (unsigned long long)p->workspace_bytes);
  uint8_t* dense = static_cast<uint8_t*>(p->workspace);
  cudaError_t e = cudaMemsetAsync(dense, 0, need, st);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(blockmask)");
  e = launch_expand_blockmask(p->blockmask, p->blockmask_nrow, p->blockmask_ncol, dense, st);
  if (e != cudaSuccess) return cuda_fail(e, "expand_blockmask launch");
  bm->dense = dense;
  bm->nrow = p->blockmask_nrow;
  bm->ncol = p->blockmask_ncol;
  return FA_OK;
}

int run_fwd(const fa_params* p, void* stream, bool block) {
  g_last_error[0] = 0;
  int rc = validate_common(p, /*bwd=*/false, block);
  if (rc != FA_OK) return rc;
  rc = check_device();
  if (rc != FA_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  BlockMaskView bm{nullptr, 0, 0};
  if (block) {
    rc = prepare_blockmask(p, st, &bm);
    if (rc != FA_OK) return rc;
  }
  if (p->impl == FA_IMPL_SIMT) {
    cudaError_t e = launch_simt_fwd(*p, bm, /*only_sdmask=*/false, st);
    if (e != cudaSuccess) return cuda_fail(e, "simt forward launch");
    return FA_OK;
  }
  const char* msg = nullptr;
  cudaError_t e = launch_fwd_sm100(*p, bm, st, &msg);
  if (e != cudaSuccess) return msg ? fail(FA_ERR_INVALID_ARGUMENT, "%s", msg) : cuda_fail(e, "tcgen05 forward launch");
  if (p->s_dmask != nullptr) {
    // debug output: +-P from the final lse and the position-keyed dropout stream
    e = launch_simt_fwd(*p, bm, /*only_sdmask=*/true, st);
    if (e != cudaSuccess) return cuda_fail(e, "S_dmask kernel launch");
  }
  return FA_OK;
}

int run_bwd(const fa_params* p, void* stream, bool block) {
  g_last_error[0] = 0;
  int rc = validate_common(p, /*bwd=*/true, block);
  if (rc != FA_OK) return rc;
  rc = check_device();
  if (rc != FA_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  BlockMaskView bm{nullptr, 0, 0};
  if (block) {
    rc = prepare_blockmask(p, st, &bm);
    if (rc != FA_OK) return rc;
  }
  if (p->impl == FA_IMPL_SIMT) {
    cudaError_t e = launch_simt_bwd(*p, bm, st);
    if (e != cudaSuccess) return cuda_fail(e, "simt backward launch");
    return FA_OK;
  }
  const uint64_t need = fa_workspace_bytes(p, block ? FA_OP_BWD_BLOCK : FA_OP_BWD);
  if (need > 0 && (p->workspace == nullptr || p->workspace_bytes < need))
    return fail(FA_ERR_WORKSPACE, "workspace too small: need %llu bytes, have %llu", (unsigned long long)need,
                (unsigned long long)p->workspace_bytes);
  const char* msg = nullptr;
  void* bwd_ws = static_cast<uint8_t*>(p->workspace) + (block ? blockmask_bytes(p) : 0);
  cudaError_t e = launch_bwd_sm100(*p, bm, bwd_ws, st, &msg);
  if (e != cudaSuccess) return msg ? fail(FA_ERR_INVALID_ARGUMENT, "%s", msg) : cuda_fail(e, "tcgen05 backward launch");
  return FA_OK;
}

}  // namespace

// ---- optional per-kernel device timing (benchmarks): cudaEvents recorded on the launching stream around the
// main forward / backward kernels while enabled; fa_timing_collect() sums and releases them. ----
struct TimedSpan { cudaEvent_t a, b; int kind; };
static std::mutex g_timing_mu;
static std::vector<TimedSpan> g_spans;
static std::atomic<int> g_timing_enabled{0};
bool timing_enabled() { return g_timing_enabled.load(std::memory_order_relaxed) != 0; }
void* timing_begin(cudaStream_t st) {
  if (!timing_enabled()) return nullptr;
  TimedSpan* sp = new TimedSpan;
  cudaEventCreate(&sp->a);
  cudaEventCreate(&sp->b);
  cudaEventRecord(sp->a, st);
  return sp;
}
void timing_end(void* h, int kind, cudaStream_t st) {
  if (h == nullptr) return;
  TimedSpan* sp = static_cast<TimedSpan*>(h);
  sp->kind = kind;
  cudaEventRecord(sp->b, st);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  g_spans.push_back(*sp);
  delete sp;
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
unsigned long long launches_so_far() { return g_launches.load(std::memory_order_relaxed); }

cudaError_t launch_expand_blockmask(const int32_t* converted, int nrow, int ncol, uint8_t* dense, cudaStream_t stream) {
  const int n = nrow * ncol;
  expand_blockmask_kernel<<<(n + 255) / 256, 256, 0, stream>>>(converted, nrow, ncol, dense);
  count_launch();
  return cudaGetLastError();
}

}  // namespace fa

// ------------------------------------------------------------------------------------------------
// exported C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {
#pragma GCC visibility push(default)

int fa_abi_version(void) { return FA_B200_ABI_VERSION; }
uint64_t fa_sizeof_params(void) { return sizeof(fa_params); }
const char* fa_last_error(void) { return fa::g_last_error; }
uint64_t fa_launch_count(void) { return fa::launches_so_far(); }

uint64_t fa_workspace_bytes(const fa_params* p, int op) {
  if (p == nullptr) return 0;
  uint64_t bytes = 0;
  if (op == FA_OP_FWD_BLOCK || op == FA_OP_BWD_BLOCK) bytes += fa::blockmask_bytes(p);
  if ((op == FA_OP_BWD || op == FA_OP_BWD_BLOCK) && p->impl != FA_IMPL_SIMT) bytes += fa::bwd_sm100_workspace_bytes(*p);
  return bytes;
}

uint64_t fa_workspace_zero_range(const fa_params* p, int op, uint64_t* begin) {
  uint64_t off = 0, len = 0;
  if (p != nullptr) {
    if (op == FA_OP_FWD_BLOCK || op == FA_OP_BWD_BLOCK) off = fa::blockmask_bytes(p);
    if ((op == FA_OP_BWD || op == FA_OP_BWD_BLOCK) && p->impl != FA_IMPL_SIMT) len = fa::bwd_sm100_workspace_zero_bytes(*p);
  }
  if (begin != nullptr) *begin = off;
  return len;
}

int fa_philox_advance(uint64_t* state, uint64_t* out, uint64_t increment, void* stream) {
  fa::g_last_error[0] = 0;
  if (state == nullptr || out == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_philox_advance: NULL pointer");
  fa::philox_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(state, out, increment);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fa::cuda_fail(e, "philox_advance launch");
  return FA_OK;
}

// The dropout stream is keyed by position, so one call consumes one "slot" of the generator; the increment is
// kept a multiple of 4 like ATen's philox_cuda_state (fmha_api.cpp:228-234).
uint64_t fa_philox_increment(const fa_params* p) { return (p != nullptr && p->p_dropout > 0.f) ? 4 : 0; }

void fa_timing_enable(int on) { fa::g_timing_enabled.store(on ? 1 : 0); }
int fa_timing_collect(double* fwd_ms, uint64_t* fwd_launches, double* bwd_ms, uint64_t* bwd_launches) {
  std::lock_guard<std::mutex> lk(fa::g_timing_mu);
  double ms[2] = {0.0, 0.0};
  uint64_t n[2] = {0, 0};
  int rc = FA_OK;
  for (auto& sp : fa::g_spans) {
    float t = 0.f;
    cudaError_t e = cudaEventSynchronize(sp.b);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&t, sp.a, sp.b);
    if (e != cudaSuccess) rc = fa::cuda_fail(e, "fa_timing_collect");
    else { ms[sp.kind & 1] += t; n[sp.kind & 1] += 1; }
    cudaEventDestroy(sp.a);
    cudaEventDestroy(sp.b);
  }
  fa::g_spans.clear();
  if (fwd_ms) *fwd_ms = ms[0];
  if (fwd_launches) *fwd_launches = n[0];
  if (bwd_ms) *bwd_ms = ms[1];
  if (bwd_launches) *bwd_launches = n[1];
  return rc;
}

int fa_fwd(const fa_params* p, void* stream) { return fa::run_fwd(p, stream, false); }
int fa_bwd(const fa_params* p, void* stream) { return fa::run_bwd(p, stream, false); }
// ---- fa_bwd_static_scratch: the backward with its scratch in a static device array (no allocation anywhere) ----
// The array lives in this translation unit.  In HOST code its name is only a host-side stand-in for the symbol, not a
// device address: the device address is obtained once per device, with that device current, by cudaGetSymbolAddress
// (below) and that pointer is what the kernels receive as their workspace.
static __device__ __align__(256) unsigned char g_static_scratch[FA_STATIC_SCRATCH_BYTES];

namespace fa {
namespace {
// -> device address of g_static_scratch on the CURRENT device (cached per device ordinal), or nullptr + *err
void* static_scratch_ptr(cudaError_t* err) {
  static std::atomic<void*> cached[64];   // nullptr = not resolved yet; concurrent first calls just resolve twice
  *err = cudaSuccess;
  int dev = 0;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  void* ptr = (dev >= 0 && dev < 64) ? cached[dev].load(std::memory_order_acquire) : nullptr;
  if (ptr == nullptr) {
    *err = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (*err != cudaSuccess) return nullptr;
    if (dev >= 0 && dev < 64) cached[dev].store(ptr, std::memory_order_release);
  }
  return ptr;
}
}  // namespace
}  // namespace fa

int fa_bwd_static_scratch(const fa_params* p, void* stream) {
  fa::g_last_error[0] = 0;
  if (p == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "params is NULL");
  if (p->impl == FA_IMPL_SIMT)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_bwd_static_scratch runs the tcgen05 kernels only (impl must not be FA_IMPL_SIMT)");
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  fa_params q = *p;
  q.workspace = nullptr;
  q.workspace_bytes = 0;
  const uint64_t need = fa_workspace_bytes(&q, FA_OP_BWD);
  if (need > FA_STATIC_SCRATCH_BYTES)
    return fa::fail(FA_ERR_WORKSPACE, "fa_bwd_static_scratch: these sizes need %llu bytes of scratch, the static array holds %llu "
                    "(FA_STATIC_SCRATCH_BYTES); use fa_bwd with a caller-owned workspace",
                    (unsigned long long)need, (unsigned long long)FA_STATIC_SCRATCH_BYTES);
  cudaError_t e;
  void* ptr = fa::static_scratch_ptr(&e);
  if (ptr == nullptr) return fa::cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
  q.workspace = ptr;
  q.workspace_bytes = FA_STATIC_SCRATCH_BYTES;
  // the array keeps the previous call's contents (possibly of other sizes): without this flag the backward clears, on
  // `stream`, the range that has to start at zero (cudaMemsetAsync -- not an allocation)
  q.flags &= ~static_cast<uint32_t>(FA_FLAG_WS_ZEROED);
  return fa::run_bwd(&q, stream, false);
}

int fa_bwd_dq_finalize(const fa_params* p, void* stream) {
  fa::g_last_error[0] = 0;
  if (p == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "params is NULL");
  if (p->b <= 0 || p->h <= 0 || p->d <= 0 || p->d % 8 != 0 || p->d > 128 || p->total_q < 0 || p->max_seqlen_q < 0)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_bwd_dq_finalize: bad sizes");
  if (p->cu_seqlens_q == nullptr || p->cu_seqlens_k == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_bwd_dq_finalize: cu_seqlens are NULL");
  if (p->total_q > 0 && (p->dq == nullptr || !fa::aligned16(p->dq) || p->dq_row_stride % 8 != 0 || p->dq_head_stride % 8 != 0))
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_bwd_dq_finalize: dq must be non-NULL, 16-byte aligned, strides multiples of 8");
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  const uint64_t need = fa::bwd_sm100_workspace_bytes(*p);
  if (p->workspace == nullptr || p->workspace_bytes < need)
    return fa::fail(FA_ERR_WORKSPACE, "workspace too small: need %llu bytes, have %llu", (unsigned long long)need,
                    (unsigned long long)p->workspace_bytes);
  cudaError_t e = fa::launch_bwd_dq_finalize(*p, p->workspace, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "dq finalize launch");
}

int fa_fwd_block(const fa_params* p, void* stream) { return fa::run_fwd(p, stream, true); }
int fa_bwd_block(const fa_params* p, void* stream) { return fa::run_bwd(p, stream, true); }

int fa_merge_partials(float* o_acc, float* lse_acc, const void* o_blk, const float* lse_blk, int32_t b, int32_t h, int32_t d,
                      int32_t seqlen, int64_t batch_rows, int32_t lse_acc_stride, int32_t lse_blk_stride, int32_t is_bf16,
                      int32_t first, void* stream) {
  fa::g_last_error[0] = 0;
  if (b <= 0 || h <= 0 || seqlen < 0 || d <= 0 || d % 8 != 0 || d > 128)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_merge_partials: bad sizes (b %d, h %d, d %d, seqlen %d)", b, h, d, seqlen);
  if (seqlen == 0) return FA_OK;
  if (!o_acc || !lse_acc || !o_blk || !lse_blk) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_merge_partials: NULL pointer");
  if (lse_acc_stride < seqlen || lse_blk_stride < seqlen || batch_rows < seqlen)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_merge_partials: lse stride or batch_rows < seqlen");
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  int d8p = 1;
  while (d8p < d / 8) d8p *= 2;
  const int64_t total = (int64_t)b * seqlen * h * d8p;
  const unsigned grid = static_cast<unsigned>((total + 255) / 256);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (is_bf16) fa::merge_partials_kernel<true><<<grid, 256, 0, st>>>(o_acc, lse_acc, o_blk, lse_blk, b, h, d, seqlen, batch_rows, lse_acc_stride, lse_blk_stride, d8p, first);
  else fa::merge_partials_kernel<false><<<grid, 256, 0, st>>>(o_acc, lse_acc, o_blk, lse_blk, b, h, d, seqlen, batch_rows, lse_acc_stride, lse_blk_stride, d8p, first);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fa::cuda_fail(e, "merge_partials launch");
  return FA_OK;
}

int fa_accumulate_f32(float* dst, const void* src, int64_t n_batches, int64_t run_elems, int64_t dst_batch_stride,
                      int64_t src_batch_stride, int32_t is_bf16, void* stream) {
  fa::g_last_error[0] = 0;
  if (n_batches < 0 || run_elems < 0 || run_elems % 8 != 0 || dst_batch_stride % 4 != 0 || src_batch_stride % 8 != 0)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_accumulate_f32: sizes and strides must be multiples of 8 elements");
  if (n_batches == 0 || run_elems == 0) return FA_OK;
  if (!dst || !src || !fa::aligned16(dst) || !fa::aligned16(src))
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_accumulate_f32: NULL or misaligned pointer");
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  const int64_t want = (n_batches * (run_elems / 8) + 255) / 256;
  const unsigned grid = static_cast<unsigned>(want < 148 * 16 ? want : 148 * 16);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (is_bf16) fa::accumulate_f32_kernel<true><<<grid, 256, 0, st>>>(dst, src, n_batches, run_elems, dst_batch_stride, src_batch_stride);
  else fa::accumulate_f32_kernel<false><<<grid, 256, 0, st>>>(dst, src, n_batches, run_elems, dst_batch_stride, src_batch_stride);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fa::cuda_fail(e, "accumulate_f32 launch");
  return FA_OK;
}

int fa_flag_write(uint64_t* flag, uint64_t value, void* stream) {
  fa::g_last_error[0] = 0;
  if (flag == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_flag_write: NULL flag");
  fa::flag_write_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, value);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "flag_write launch");
}

int fa_flag_wait_geq(const uint64_t* flag, uint64_t value, uint64_t timeout_ms, void* stream) {
  fa::g_last_error[0] = 0;
  if (flag == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_flag_wait_geq: NULL flag");
  if (timeout_ms == 0) timeout_ms = 20000;
  fa::flag_wait_geq_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, value, timeout_ms * 1000000ull);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "flag_wait_geq launch");
}

// CUDA IPC for the peer-memory ring.  The handle names the whole cudaMalloc allocation that contains `ptr` (a framework's
// caching allocator hands out sub-ranges), so the offset of `ptr` inside it travels with the handle.
int fa_ipc_export(const void* ptr, uint8_t handle[64], uint64_t* offset) {
  fa::g_last_error[0] = 0;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  if (ptr == nullptr || handle == nullptr || offset == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_ipc_export: NULL pointer");
  typedef CUresult (*GetRange)(CUdeviceptr*, size_t*, CUdeviceptr);
  static GetRange get_range = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
    return reinterpret_cast<GetRange>(fn);
  }();
  if (get_range == nullptr) return fa::fail(FA_ERR_CUDA, "fa_ipc_export: cuMemGetAddressRange is not available");
  CUdeviceptr base = 0;
  size_t size = 0;
  if (get_range(&base, &size, reinterpret_cast<CUdeviceptr>(ptr)) != CUDA_SUCCESS)
    return fa::fail(FA_ERR_CUDA, "fa_ipc_export: cuMemGetAddressRange failed (not a device allocation?)");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(base));
  if (e != cudaSuccess) return fa::cuda_fail(e, "cudaIpcGetMemHandle");
  memcpy(handle, &h, 64);
  *offset = reinterpret_cast<CUdeviceptr>(ptr) - base;
  return FA_OK;
}

int fa_ipc_open(const uint8_t handle[64], void** base) {
  fa::g_last_error[0] = 0;
  if (handle == nullptr || base == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_ipc_open: NULL pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  // opened with the CALLER's device current: the mapping belongs to this device's context and peer access to the
  // exporting GPU is enabled along with it, so this device's kernels and copies can use the addresses directly
  cudaError_t e = cudaIpcOpenMemHandle(base, h, cudaIpcMemLazyEnablePeerAccess);
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "cudaIpcOpenMemHandle");
}

int fa_ipc_close(void* base) {
  fa::g_last_error[0] = 0;
  if (base == nullptr) return FA_OK;
  cudaError_t e = cudaIpcCloseMemHandle(base);
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "cudaIpcCloseMemHandle");
}

int fa_enable_peer_access(int32_t peer_device) {
  fa::g_last_error[0] = 0;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fa::cuda_fail(e, "cudaGetDevice");
  if (peer_device == dev) return FA_OK;
  int can = 0;
  e = cudaDeviceCanAccessPeer(&can, dev, peer_device);
  if (e != cudaSuccess) return fa::cuda_fail(e, "cudaDeviceCanAccessPeer");
  if (!can) return fa::fail(FA_ERR_UNSUPPORTED_DEVICE, "device %d cannot access device %d as a peer (no NVLink / P2P path)", dev, peer_device);
  e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return FA_OK; }   // clear the sticky-less error state
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "cudaDeviceEnablePeerAccess");
}

int fa_peer_copy(void* dst, const void* src, uint64_t bytes, void* stream) {
  fa::g_last_error[0] = 0;
  if (bytes == 0) return FA_OK;
  if (dst == nullptr || src == nullptr) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_peer_copy: NULL pointer");
  // unified addressing: the driver sees that src lives on a peer device and uses a copy engine over NVLink
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "cudaMemcpyAsync (peer)");
}

int fa_gather_rows(void* dst, const void* src, const int64_t* indices, int64_t n_indices, int64_t row_bytes, void* stream) {
  fa::g_last_error[0] = 0;
  return fa::move_rows<true>(dst, src, indices, n_indices, row_bytes, stream);
}
int fa_scatter_rows(void* dst, const void* src, const int64_t* indices, int64_t n_indices, int64_t row_bytes, void* stream) {
  fa::g_last_error[0] = 0;
  return fa::move_rows<false>(dst, src, indices, n_indices, row_bytes, stream);
}

int fa_scatter_add_rows(void* dst, const void* src, const int64_t* indices, int64_t n_indices, int64_t row_bytes, int32_t kind,
                        void* stream) {
  fa::g_last_error[0] = 0;
  if (n_indices < 0 || row_bytes < 0 || kind < 0 || kind > 2) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_scatter_add_rows: bad size or kind");
  if (n_indices == 0 || row_bytes == 0) return FA_OK;
  if (!dst || !src || !indices || !fa::aligned16(dst) || !fa::aligned16(src) || row_bytes % 16 != 0)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_scatter_add_rows: pointers must be 16-byte aligned and rows a multiple of 16 bytes (got %lld)",
                    (long long)row_bytes);
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  const int64_t vecs = row_bytes / 16;
  int threads = 32;
  while (threads < 256 && threads < vecs) threads *= 2;
  const unsigned grid = static_cast<unsigned>(n_indices < 148 * 16 ? n_indices : 148 * 16);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (kind == 0) fa::scatter_add_rows_kernel<0><<<grid, threads, 0, st>>>(dst, src, indices, n_indices, vecs);
  else if (kind == 1) fa::scatter_add_rows_kernel<1><<<grid, threads, 0, st>>>(dst, src, indices, n_indices, vecs);
  else fa::scatter_add_rows_kernel<2><<<grid, threads, 0, st>>>(dst, src, indices, n_indices, vecs);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "scatter_add_rows launch");
}

int fa_varlen_index(const uint8_t* mask, int32_t b, int32_t s, int64_t* indices, int32_t* cu_seqlens, int32_t* stats, void* stream) {
  fa::g_last_error[0] = 0;
  if (b <= 0 || s <= 0) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_varlen_index: batch and seqlen must be positive (got %d x %d)", b, s);
  if (!mask || !indices || !cu_seqlens || !stats) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_varlen_index: NULL pointer");
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(indices, 0xFF, (size_t)b * s * sizeof(int64_t), st);     // -1 = "no token": entries past the total
  if (e != cudaSuccess) return fa::cuda_fail(e, "cudaMemsetAsync(indices)");
  fa::varlen_count_kernel<<<b, 256, 0, st>>>(mask, s, cu_seqlens);
  fa::varlen_scan_kernel<<<1, 1024, 0, st>>>(cu_seqlens, b, stats);
  fa::varlen_indices_kernel<<<b, 256, 0, st>>>(mask, s, cu_seqlens, indices);
  fa::count_launch(); fa::count_launch(); fa::count_launch();
  e = cudaGetLastError();
  return e == cudaSuccess ? FA_OK : fa::cuda_fail(e, "varlen_index launch");
}

int fa_rotary_qk(const void* in, void* out, const float* cos_t, const float* sin_t, int64_t n_rows, int32_t seqlen, int32_t n_parts,
                 int32_t n_rot, int32_t h, int32_t d, int64_t in_row_stride, int64_t out_row_stride, int32_t conjugate,
                 int32_t is_bf16, void* stream) {
  fa::g_last_error[0] = 0;
  if (n_rows < 0 || seqlen <= 0 || n_parts <= 0 || n_rot < 0 || n_rot > n_parts || h <= 0 || d <= 0 || d % 8 != 0)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_rotary_qk: bad sizes (rows %lld, seqlen %d, parts %d/%d, h %d, d %d)",
                    (long long)n_rows, seqlen, n_rot, n_parts, h, d);
  if (n_rows == 0) return FA_OK;
  if (!in || !out || !cos_t || !sin_t) return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_rotary_qk: NULL pointer");
  const int64_t row_elems = (int64_t)n_parts * h * d;
  if (in_row_stride < row_elems || out_row_stride < row_elems || in_row_stride % 8 != 0 || out_row_stride % 8 != 0)
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_rotary_qk: row strides must be >= parts*h*d and multiples of 8 elements");
  if (!fa::aligned16(in) || !fa::aligned16(out) || !fa::aligned16(cos_t) || !fa::aligned16(sin_t))
    return fa::fail(FA_ERR_INVALID_ARGUMENT, "fa_rotary_qk: pointers must be 16-byte aligned");
  int rc = fa::check_device();
  if (rc != FA_OK) return rc;
  // in place: only the rotated parts are touched
  const int parts = (in == out) ? n_rot : n_parts;
  if (parts == 0) return FA_OK;
  const int64_t total = n_rows * parts * h * (d / 8);
  const int64_t want = (total + 255) / 256;
  const unsigned grid = static_cast<unsigned>(want < 148 * 16 ? want : 148 * 16);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float sign = conjugate ? -1.f : 1.f;
  if (is_bf16) fa::rotary_qk_kernel<true><<<grid, 256, 0, st>>>(in, out, cos_t, sin_t, n_rows, seqlen, parts, n_rot, h, d, in_row_stride, out_row_stride, sign);
  else fa::rotary_qk_kernel<false><<<grid, 256, 0, st>>>(in, out, cos_t, sin_t, n_rows, seqlen, parts, n_rot, h, d, in_row_stride, out_row_stride, sign);
  fa::count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fa::cuda_fail(e, "rotary_qk launch");
  return FA_OK;
}

void fa_dropout_keep8_host(uint64_t seed, uint64_t offset, uint32_t bh, uint32_t qrow, uint32_t kgroup, float p_dropout,
                           uint8_t* keep) {
  const fa::PhiloxKey pk = fa::make_philox_key(seed, offset);
  const uint4 r = fa::dropout_rand8(pk, pk.c3, bh, qrow, kgroup);
  const uint32_t thr = fa::keep_threshold_u16(p_dropout);
  for (int e = 0; e < 8; ++e) keep[e] = fa::rand8_get(r, e) <= thr ? 1 : 0;
}

#pragma GCC visibility pop
}  // extern "C"
