/* fa_b200.h -- C ABI of libfa_b200.so, the B200 (sm_100a) replacement for the reference's
 * `flash_attn_cuda` native extension on the attention hot path.
 *
 * Every entry point below replaces one native call the reference's Python layer binds
 * (paths relative to the reference repo):
 *
 *   fa_fwd        <- flash_attn_cuda.fwd        csrc/flash_attn/fmha_api.cpp:112-242 (mha_fwd),
 *                                               called at flash_attn/flash_attn_interface.py:19-22
 *   fa_bwd        <- flash_attn_cuda.bwd        called at flash_attn/flash_attn_interface.py:31-33
 *                                               (native source absent from the reference snapshot)
 *   fa_fwd_block  <- flash_attn_cuda.fwd_block  called at flash_attn/flash_blocksparse_attn_interface.py:45-47
 *   fa_bwd_block  <- flash_attn_cuda.bwd_block  called at flash_attn/flash_blocksparse_attn_interface.py:56-58
 *   fa_rotary_qk  <- apply_rotary_pos_emb       flash_attn/rotary.py:31-42 as used by FlashMHA (flash_attention.py:105-110)
 *   fa_gather_rows / fa_scatter_rows            the torch.gather / scatter_ / index_put the reference issues in
 *                                               flash_attn/bert_padding.py:21-22, 30-35, 48-53, 59
 *
 * Conventions: plain pointers and sizes only (no torch types); all data pointers are DEVICE pointers;
 * strides are in ELEMENTS; the library never allocates device memory and never synchronises the host --
 * work is enqueued on `stream` (a cudaStream_t passed as void*).  Every function returns 0 on success and
 * a nonzero FA_ERR_* code otherwise; fa_last_error() gives the message for the calling thread.
 * Argument errors correspond to the TORCH_CHECKs of fmha_api.cpp:127-170.
 */
#ifndef FA_B200_H_
#define FA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FA_B200_ABI_VERSION 2

enum {
  FA_OK = 0,
  FA_ERR_INVALID_ARGUMENT = 1, /* shape / dtype / stride / alignment violations */
  FA_ERR_UNSUPPORTED_DEVICE = 2, /* device is not compute capability 10.x */
  FA_ERR_CUDA = 3,             /* a CUDA runtime / driver call failed */
  FA_ERR_WORKSPACE = 4         /* workspace missing or too small */
};

/* kernel selection (fa_params.impl) */
enum {
  FA_IMPL_AUTO = 0,   /* tcgen05/TMA kernels (the product path) */
  FA_IMPL_TCGEN05 = 1,
  FA_IMPL_SIMT = 2    /* CUDA-core cross-check kernels: bring-up / debugging only */
};

/* Mirrors Qkv_params + FMHA_fprop_params of the reference (csrc/flash_attn/src/fmha.h:50-124),
 * with 64-bit strides (the reference's 32-bit strides limit tensors to 2 GB, fmha.h:59-60). */
typedef struct fa_params {
  /* --- device pointers ------------------------------------------------------------------- */
  const void* q;            /* (total_q, h, d) */
  const void* k;            /* (total_k, h, d) */
  const void* v;            /* (total_k, h, d) */
  void* o;                  /* (total_q, h, d): fwd output, bwd input */
  const void* dout;         /* bwd: (total_q, h, d) */
  void* dq;                 /* bwd outputs, caller-allocated, fully overwritten */
  void* dk;
  void* dv;
  float* softmax_lse;       /* (b, h, seqlen_q_rounded) fp32, natural log; fwd output, bwd input */
  float* softmax_d;         /* bwd output (b, h, seqlen_q_rounded) fp32 = rowsum(dO * O) */
  void* s_dmask;            /* optional debug output (b, h, seqlen_q_rounded, seqlen_k_rounded), same dtype as q:
                               +P kept / -P dropped, P = exp(scale*s - lse)  (NULL = not requested) */
  const int32_t* cu_seqlens_q; /* (b + 1) */
  const int32_t* cu_seqlens_k; /* (b + 1) */
  const int32_t* blockmask; /* blocksparse only: converted mask (ncol, nrow) int32, see
                               flash_blocksparse_attn_interface.py:8-40; NULL for dense */
  void* workspace;          /* scratch, size >= fa_workspace_bytes() */
  uint64_t workspace_bytes;
  /* --- optional device pointers (NULL = feature off) --------------------------------------- */
  const int32_t* seqused_q; /* (b): rows of sequence i actually used, starting at cu_seqlens_q[i]; when given,
                               cu_seqlens_q[i+1] is not read (sequences may leave gaps).  Lets one launch address
                               sub-ranges of a packed batch (the K/V ring, flash_attn/ring_attention.py) */
  const int32_t* seqused_k; /* (b): same for keys */
  const float* alibi_slopes; /* (h) fp32: additive attention bias -slope[head] * |i - j| on the scaled scores
                               (i, j = query / key index inside the sequence); gradient-free (README.md:57 roadmap) */
  const uint64_t* philox_offset_ptr; /* device counter added (mod 2^32) to the low word of philox_offset when the
                               kernel runs: a captured CUDA graph draws a fresh dropout mask per replay, like the
                               reference's PhiloxCudaState (csrc/flash_attn/src/fmha.h:39, fmha_fprop_kernel_1xN.h:697) */

  /* --- strides, in elements (last dim must be contiguous) --------------------------------- */
  int64_t q_row_stride, q_head_stride;
  int64_t k_row_stride, k_head_stride;
  int64_t v_row_stride, v_head_stride;
  int64_t o_row_stride, o_head_stride;
  int64_t do_row_stride, do_head_stride;
  int64_t dq_row_stride, dq_head_stride;
  int64_t dk_row_stride, dk_head_stride;
  int64_t dv_row_stride, dv_head_stride;

  /* --- sizes ------------------------------------------------------------------------------ */
  int32_t b, h, d;
  int32_t total_q, total_k;
  int32_t max_seqlen_q, max_seqlen_k;          /* caller's upper bounds */
  int32_t seqlen_q_rounded, seqlen_k_rounded;  /* row / col extents of softmax_lse, softmax_d, s_dmask */
  int32_t blockmask_nrow, blockmask_ncol;      /* blocksparse: nrow = S/16 blocks, ncol = S/256 blocks */

  /* --- scalars ---------------------------------------------------------------------------- */
  float softmax_scale;
  float p_dropout;          /* probability of DROPPING, in [0, 1) */
  int32_t is_causal;        /* top-left aligned: key j visible to query i iff j <= i */
  int32_t is_bf16;          /* 0 = fp16, 1 = bf16 */
  int32_t impl;             /* FA_IMPL_* */
  uint32_t flags;           /* FA_FLAG_* */
  uint64_t philox_seed;     /* dropout RNG stream (position keyed: identical in fwd, bwd and s_dmask) */
  uint64_t philox_offset;
} fa_params;

enum { FA_OP_FWD = 0, FA_OP_BWD = 1, FA_OP_FWD_BLOCK = 2, FA_OP_BWD_BLOCK = 3 };

/* fa_params.flags */
enum {
  /* backward: the caller guarantees that the workspace bytes named by fa_workspace_zero_range() are zero on entry
   * (e.g. a buffer kept between calls on the same stream).  Every backward leaves that range zero again when it
   * completes, so a persistent workspace needs to be cleared exactly once; without the flag the library clears it
   * itself (one cudaMemsetAsync per call). */
  FA_FLAG_WS_ZEROED = 1u,
  /* backward: do not run the final pass that converts the fp32 dQ accumulator -- dQ stays in the workspace as unscaled
   * partial sums, p->dq is not written (may be NULL) and the workspace's zero range is NOT zero on return.  Further
   * fa_bwd calls on the SAME workspace with the same total_q / b / h / d keep adding to it (the K/V ring: one call per
   * ring step, each with other keys); fa_bwd_dq_finalize() then produces dq and leaves the accumulator zero again.
   * The caller owns that workspace for the whole sequence of calls. */
  FA_FLAG_KEEP_DQ_ACCUM = 2u,
  /* backward (fa_bwd, tcgen05 kernels, no block mask): bitwise reproducible dQ.  By default the key blocks of a sequence add
   * their dQ parts to a query tile with fp32 reductions in whatever order they get there, so dQ varies in the last bits from
   * run to run (dK and dV never do); with this flag they take turns in a fixed order.  The main backward kernel then takes
   * 1.2x (d = 64 causal with dropout) to 1.7x (d = 128 causal) and 2.6x (d = 128 non-causal) as long: every hand-over waits
   * for the reductions to complete.
   * Not combinable with FA_FLAG_KEEP_DQ_ACCUM.  No counterpart in the reference's interface: its kernels walk the key blocks
   * of a (batch, head) pair inside one CTA (forward: fmha_fprop_kernel_1xN.h:686-717; the backward sources are not in the
   * snapshot), an ordered accumulation by construction; this flag is how to get an ordered dQ from this library. */
  FA_FLAG_DETERMINISTIC = 4u
};

int fa_abi_version(void);
uint64_t fa_sizeof_params(void);   /* sizeof(fa_params): lets a foreign-language binding verify its struct mirror */
const char* fa_last_error(void);
/* number of CUDA kernels this library has launched in this process (monotonic; for benchmarks' launch accounting) */
uint64_t fa_launch_count(void);

/* Benchmark support: while enabled, the library brackets its main forward / backward kernels with cudaEvents on the
 * launching stream; fa_timing_collect() waits for them, returns the summed device time (ms) and launch counts since
 * the last collect, and releases the events.  Off by default (no events are created). */
void fa_timing_enable(int on);
int fa_timing_collect(double* fwd_ms, uint64_t* fwd_launches, double* bwd_ms, uint64_t* bwd_launches);

/* bytes of scratch `op` needs for these sizes (0 if none). Pure host arithmetic.  The backward's scratch is
 * O(total_q * h * d): statistics and the fp32 dQ accumulator are indexed by packed row, not by batch x max_seqlen. */
uint64_t fa_workspace_bytes(const fa_params* p, int op);
/* the part of the workspace that must be zero on entry when FA_FLAG_WS_ZEROED is set: bytes [*begin, *begin + return
 * value) (return value 0: the op has no such region).  The rest of the workspace needs no initialisation. */
uint64_t fa_workspace_zero_range(const fa_params* p, int op, uint64_t* begin);
/* Graph-capturable dropout: out[0] = state[0]; state[0] += increment -- one tiny kernel on `stream`.  A captured
 * forward passes `out` as philox_offset_ptr (and saves it for its backward); every replay then draws a new mask. */
int fa_philox_advance(uint64_t* state, uint64_t* out, uint64_t increment, void* stream);
/* how far the caller must advance its Philox offset after a call with dropout (fmha_api.cpp:228-234). */
uint64_t fa_philox_increment(const fa_params* p);

int fa_fwd(const fa_params* p, void* stream);
int fa_bwd(const fa_params* p, void* stream);
int fa_fwd_block(const fa_params* p, void* stream);
/* completes a sequence of FA_FLAG_KEEP_DQ_ACCUM backward calls: dq = softmax_scale * accumulator, 16 bit, for the query
 * layout of p (cu_seqlens_q / seqused_q / max_seqlen_q / total_q / b / h / d, p->dq and its strides, p->workspace); the
 * accumulator is zero afterwards.  Other fields of p are not used. */
int fa_bwd_dq_finalize(const fa_params* p, void* stream);
int fa_bwd_block(const fa_params* p, void* stream);

/* fa_bwd for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same contract as fa_bwd, except that p->workspace / p->workspace_bytes are ignored: the scratch (fp32 dQ accumulator,
 * tile counters, per-row statistics) is a file-scope static __device__ array of FA_STATIC_SCRATCH_BYTES inside the library
 * (every GPU that loads the library holds that much until the process exits).  The call allocates nothing -- no
 * cudaMalloc / cudaFree, no allocating library routine -- and fails with FA_ERR_WORKSPACE when
 * fa_workspace_bytes(p, FA_OP_BWD) exceeds FA_STATIC_SCRATCH_BYTES (fa_last_error() states both numbers).
 * All calls on a device share the one array, which keeps the previous call's contents: the part that must start at zero
 * is cleared on `stream` by every call.  Calls that overlap on the GPU (different streams, or graphs replayed
 * concurrently) RACE on it -- serialise them, or use fa_bwd with a caller-owned workspace. */
#define FA_STATIC_SCRATCH_BYTES (32u << 20)
int fa_bwd_static_scratch(const fa_params* p, void* stream);

/* Varlen pack / unpack (bert_padding.py).  Rows are `row_bytes` long (multiple of 4 bytes).
 * gather : dst[i, :] = src[indices[i], :]             for i in [0, n_indices); a negative index gives a zero row
 * scatter: dst[indices[i], :] = src[i, :]             (dst rows not named by indices are left untouched; negative = skip)
 * indices are int64 (torch.nonzero output). */
int fa_gather_rows(void* dst, const void* src, const int64_t* indices, int64_t n_indices, int64_t row_bytes,
                   void* stream);
int fa_scatter_rows(void* dst, const void* src, const int64_t* indices, int64_t n_indices, int64_t row_bytes,
                    void* stream);

/* dst[indices[i], :] += src[i, :] (indices unique; negative entries are skipped) -- the backward of
 * index_first_axis_residual (flash_attn/bert_padding.py:67-96, `grad_residual.index_add(0, indices, grad_output)`).
 * kind: 0 = fp16, 1 = bf16, 2 = fp32; rows a multiple of 16 bytes, 16-byte aligned. */
int fa_scatter_add_rows(void* dst, const void* src, const int64_t* indices, int64_t n_indices, int64_t row_bytes,
                        int32_t kind, void* stream);

/* Varlen bookkeeping on the device, no host round trip (the reference's unpad_input does nonzero() and .item():
 * flash_attn/bert_padding.py:99-119).  mask: (b, s) bytes, nonzero = valid token.  Outputs:
 *   indices     (b * s) int64: flat positions of the valid tokens in order, then -1 up to the end (fa_gather_rows turns a
 *               -1 into a zero row, fa_scatter_rows / fa_scatter_add_rows skip it);
 *   cu_seqlens  (b + 1) int32;   stats[0] = longest sequence, stats[1] = number of valid tokens. */
int fa_varlen_index(const uint8_t* mask, int32_t b, int32_t s, int64_t* indices, int32_t* cu_seqlens, int32_t* stats,
                    void* stream);

/* Merge of two partial attention results over the same queries (K/V split across ring steps / GPUs): the reference's
 * inter-K-block rule (csrc/flash_attn/src/fmha_fprop_kernel_1xN.h:606-650) applied to whole blocks, in place:
 *   lse' = log(exp(lse_acc) + exp(lse_blk));  o_acc = o_acc * exp(lse_acc - lse') + o_blk * exp(lse_blk - lse');  lse_acc = lse'
 * o_acc fp32 and o_blk 16 bit: (rows, h, d) contiguous, the rows of batch entry i being i * batch_rows + t, t < seqlen
 * (batch_rows > seqlen addresses one part of every sequence); lse_acc (b, h, lse_acc_stride), lse_blk (b, h, lse_blk_stride)
 * fp32.  first != 0: initialise (o_acc = o_blk, lse_acc = lse_blk). */
int fa_merge_partials(float* o_acc, float* lse_acc, const void* o_blk, const float* lse_blk, int32_t b, int32_t h, int32_t d,
                      int32_t seqlen, int64_t batch_rows, int32_t lse_acc_stride, int32_t lse_blk_stride, int32_t is_bf16,
                      int32_t first, void* stream);

/* dst (fp32) += src (fp16 / bf16) over n_batches runs of run_elems contiguous elements (multiples of 8; strides in
 * elements): how the ring backward adds a step's dQ / dK / dV block into its fp32 accumulators. */
int fa_accumulate_f32(float* dst, const void* src, int64_t n_batches, int64_t run_elems, int64_t dst_batch_stride,
                      int64_t src_batch_stride, int32_t is_bf16, void* stream);

/* Peer-memory transport of the K/V ring (flash_attn/ring_attention.py; no counterpart in the single-GPU reference).
 * One process per GPU; every rank maps its neighbours' buffers and flag words through CUDA IPC.  Blocks move as copy-
 * engine transfers over NVLink on the caller's stream; ordering across processes goes through monotone 64-bit counters:
 *   fa_flag_write    *flag = value, published with a system-scope fence after everything already enqueued on `stream`
 *                    (flag may live in a peer GPU's memory);
 *   fa_flag_wait_geq holds `stream` until *flag >= value (a one-thread spin kernel; flag should be in LOCAL memory).  If
 *                    that does not happen within timeout_ms (0 = 20 s) the kernel traps: a lost peer surfaces as a CUDA
 *                    error on this stream, never as a hung GPU;
 *   fa_peer_copy     cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, stream) -- either pointer may be peer memory;
 *   fa_enable_peer_access  cudaDeviceEnablePeerAccess(peer_device) for the current device (idempotent): needed before this
 *                    device's kernels write a peer's flag word; FA_ERR_UNSUPPORTED_DEVICE if there is no P2P path. */
int fa_enable_peer_access(int32_t peer_device);
/* CUDA IPC: fa_ipc_export names the cudaMalloc allocation that contains `ptr` (64-byte handle) and ptr's offset in it;
 * fa_ipc_open maps such an allocation of ANOTHER process into the current device's address space (peer access to the
 * exporting GPU is enabled with it) -- open a given handle once per process; fa_ipc_close unmaps.  The exporter must
 * keep the allocation alive while it is mapped elsewhere. */
int fa_ipc_export(const void* ptr, uint8_t handle[64], uint64_t* offset);
int fa_ipc_open(const uint8_t handle[64], void** base);
int fa_ipc_close(void* base);
int fa_flag_write(uint64_t* flag, uint64_t value, void* stream);
int fa_flag_wait_geq(const uint64_t* flag, uint64_t value, uint64_t timeout_ms, void* stream);
int fa_peer_copy(void* dst, const void* src, uint64_t bytes, void* stream);

/* Rotary position embedding of the first n_rot parts (q, k) of a packed (n_rows, n_parts, h, d) 16-bit tensor in one
 * pass -- replaces the pure-PyTorch apply_rotary_pos_emb + torch.stack the reference runs before attention in FlashMHA
 * (flash_attn/rotary.py:21-42, flash_attn/flash_attention.py:105-110).  Interleaved pairs:
 *   (x0, x1) -> (x0 cos - x1 sin, x1 cos + x0 sin);  conjugate != 0 applies the transpose (the backward).
 * cos_t / sin_t: fp32 (seqlen, d/2); position of a row = row % seqlen.  Row strides in elements.  in == out rotates in
 * place; otherwise parts >= n_rot (v) are copied. */
int fa_rotary_qk(const void* in, void* out, const float* cos_t, const float* sin_t, int64_t n_rows, int32_t seqlen,
                 int32_t n_parts, int32_t n_rot, int32_t h, int32_t d, int64_t in_row_stride, int64_t out_row_stride,
                 int32_t conjugate, int32_t is_bf16, void* stream);

/* Host-only helper exposing the dropout RNG for tests: fills keep[8] (0/1) for the 8 key positions
 * 8*kgroup .. 8*kgroup+7 of query row `qrow` in (batch*h + head) = bh.  Pure CPU code. */
void fa_dropout_keep8_host(uint64_t seed, uint64_t offset, uint32_t bh, uint32_t qrow, uint32_t kgroup,
                           float p_dropout, uint8_t* keep);

#ifdef __cplusplus
}
#endif
#endif /* FA_B200_H_ */
