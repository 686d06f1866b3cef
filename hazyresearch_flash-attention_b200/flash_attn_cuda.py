"""`flash_attn_cuda` -- drop-in for the reference's pybind11 extension of the same name
(reference: csrc/flash_attn/fmha_api.cpp:244-247 exports `fwd`; the Python layer additionally calls `bwd`,
`fwd_block`, `bwd_block`: flash_attn/flash_attn_interface.py:19,31, flash_attn/flash_blocksparse_attn_interface.py:45,56).

This module is a thin ctypes binding over the C-ABI library `csrc/libfa_b200.so` (include/fa_b200.h).  It does
what `mha_fwd` does on the host (fmha_api.cpp:127-241): argument checks, output allocation, device guard,
current stream, Philox (seed, offset) from the CUDA generator -- and nothing else.  There is no fallback of
any kind: if the library is missing the import fails, and on a non-sm_100 device every call raises.
"""
import ctypes
import math
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("FA_B200_LIB") or os.path.join(_HERE, "csrc", "libfa_b200.so")   # override: debug builds only

FA_IMPL_AUTO, FA_IMPL_TCGEN05, FA_IMPL_SIMT = 0, 1, 2
FA_OP_FWD, FA_OP_BWD, FA_OP_FWD_BLOCK, FA_OP_BWD_BLOCK = 0, 1, 2, 3

FA_FLAG_WS_ZEROED, FA_FLAG_KEEP_DQ_ACCUM, FA_FLAG_DETERMINISTIC = 1, 2, 4

# fa_params.impl of every call.  Always FA_IMPL_AUTO (the tcgen05 kernels) in the product; the GPU test-suite patches
# this module attribute to FA_IMPL_SIMT to cross-check against the library's CUDA-core kernels (tests/fa_test_utils.py).
_impl = FA_IMPL_AUTO

# Bitwise reproducible dQ (FA_FLAG_DETERMINISTIC, include/fa_b200.h): None = follow torch.are_deterministic_algorithms_enabled().
# dK, dV and the forward are reproducible in either mode; the blocksparse backward and the K/V ring's accumulating calls
# (dq_workspace=) have no ordered mode and keep their fp32 reductions in arrival order.
_deterministic = None


def set_deterministic(mode):
    """True / False: dQ of `bwd` is / is not accumulated in key-block order; None: follow torch's global switch
    (torch.use_deterministic_algorithms).  Returns the previous setting."""
    global _deterministic
    if mode is not None and not isinstance(mode, bool):
        raise TypeError("set_deterministic expects True, False or None")
    prev, _deterministic = _deterministic, mode
    return prev


def _want_deterministic():
    return torch.are_deterministic_algorithms_enabled() if _deterministic is None else _deterministic


class FaParams(ctypes.Structure):
    """Mirror of `struct fa_params` in include/fa_b200.h (field order and types must match)."""
    _fields_ = (
        [(n, ctypes.c_void_p) for n in (
            "q", "k", "v", "o", "dout", "dq", "dk", "dv", "softmax_lse", "softmax_d", "s_dmask",
            "cu_seqlens_q", "cu_seqlens_k", "blockmask", "workspace")]
        + [("workspace_bytes", ctypes.c_uint64)]
        + [(n, ctypes.c_void_p) for n in ("seqused_q", "seqused_k", "alibi_slopes", "philox_offset_ptr")]
        + [(n, ctypes.c_int64) for n in (
            "q_row_stride", "q_head_stride", "k_row_stride", "k_head_stride", "v_row_stride", "v_head_stride",
            "o_row_stride", "o_head_stride", "do_row_stride", "do_head_stride", "dq_row_stride", "dq_head_stride",
            "dk_row_stride", "dk_head_stride", "dv_row_stride", "dv_head_stride")]
        + [(n, ctypes.c_int32) for n in (
            "b", "h", "d", "total_q", "total_k", "max_seqlen_q", "max_seqlen_k", "seqlen_q_rounded",
            "seqlen_k_rounded", "blockmask_nrow", "blockmask_ncol")]
        + [("softmax_scale", ctypes.c_float), ("p_dropout", ctypes.c_float)]
        + [(n, ctypes.c_int32) for n in ("is_causal", "is_bf16", "impl")] + [("flags", ctypes.c_uint32)]
        + [("philox_seed", ctypes.c_uint64), ("philox_offset", ctypes.c_uint64)]
    )


def _load_library():
    if not os.path.exists(_LIB_PATH):
        raise ImportError(
            f"{_LIB_PATH} not found: build it with `python __graft_entry__.py` (or `make -C "
            f"{os.path.dirname(_LIB_PATH)}`).  flash_attn_cuda has no fallback path.")
    lib = ctypes.CDLL(_LIB_PATH)
    lib.fa_abi_version.restype = ctypes.c_int
    lib.fa_last_error.restype = ctypes.c_char_p
    lib.fa_workspace_bytes.restype = ctypes.c_uint64
    lib.fa_workspace_bytes.argtypes = [ctypes.POINTER(FaParams), ctypes.c_int]
    lib.fa_philox_increment.restype = ctypes.c_uint64
    lib.fa_philox_increment.argtypes = [ctypes.POINTER(FaParams)]
    for name in ("fa_fwd", "fa_bwd", "fa_fwd_block", "fa_bwd_block", "fa_bwd_static_scratch", "fa_bwd_dq_finalize"):
        fn = getattr(lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = [ctypes.POINTER(FaParams), ctypes.c_void_p]
    for name in ("fa_gather_rows", "fa_scatter_rows"):
        fn = getattr(lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                       ctypes.c_void_p]
    lib.fa_workspace_zero_range.restype = ctypes.c_uint64
    lib.fa_workspace_zero_range.argtypes = [ctypes.POINTER(FaParams), ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
    lib.fa_philox_advance.restype = ctypes.c_int
    lib.fa_philox_advance.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
    lib.fa_merge_partials.restype = ctypes.c_int
    lib.fa_merge_partials.argtypes = ([ctypes.c_void_p] * 4 + [ctypes.c_int32] * 4 + [ctypes.c_int64] + [ctypes.c_int32] * 4
                                      + [ctypes.c_void_p])
    lib.fa_accumulate_f32.restype = ctypes.c_int
    lib.fa_accumulate_f32.argtypes = [ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_int64] * 4 + [ctypes.c_int32, ctypes.c_void_p]
    lib.fa_rotary_qk.restype = ctypes.c_int
    lib.fa_rotary_qk.argtypes = ([ctypes.c_void_p] * 4 + [ctypes.c_int64] + [ctypes.c_int32] * 5 + [ctypes.c_int64] * 2
                                 + [ctypes.c_int32] * 2 + [ctypes.c_void_p])
    lib.fa_flag_write.restype = ctypes.c_int
    lib.fa_flag_write.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
    lib.fa_flag_wait_geq.restype = ctypes.c_int
    lib.fa_flag_wait_geq.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p]
    lib.fa_scatter_add_rows.restype = ctypes.c_int
    lib.fa_scatter_add_rows.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                                        ctypes.c_int32, ctypes.c_void_p]
    lib.fa_varlen_index.restype = ctypes.c_int
    lib.fa_varlen_index.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p,
                                    ctypes.c_void_p, ctypes.c_void_p]
    lib.fa_ipc_export.restype = ctypes.c_int
    lib.fa_ipc_export.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64)]
    lib.fa_ipc_open.restype = ctypes.c_int
    lib.fa_ipc_open.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
    lib.fa_ipc_close.restype = ctypes.c_int
    lib.fa_ipc_close.argtypes = [ctypes.c_void_p]
    lib.fa_enable_peer_access.restype = ctypes.c_int
    lib.fa_enable_peer_access.argtypes = [ctypes.c_int32]
    lib.fa_peer_copy.restype = ctypes.c_int
    lib.fa_peer_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
    lib.fa_dropout_keep8_host.restype = None
    lib.fa_dropout_keep8_host.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32,
                                          ctypes.c_uint32, ctypes.c_float, ctypes.c_void_p]
    if lib.fa_abi_version() != 2:
        raise ImportError(f"{_LIB_PATH}: ABI version {lib.fa_abi_version()} != 2")
    lib.fa_sizeof_params.restype = ctypes.c_uint64
    lib.fa_launch_count.restype = ctypes.c_uint64
    lib.fa_timing_enable.restype = None
    lib.fa_timing_enable.argtypes = [ctypes.c_int]
    lib.fa_timing_collect.restype = ctypes.c_int
    lib.fa_timing_collect.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_uint64),
                                      ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_uint64)]
    if lib.fa_sizeof_params() != ctypes.sizeof(FaParams):
        raise ImportError(f"{_LIB_PATH}: sizeof(fa_params) = {lib.fa_sizeof_params()} but the ctypes mirror has "
                          f"{ctypes.sizeof(FaParams)} bytes")
    return lib


_lib = _load_library()


def _check(cond, msg):
    if not cond:
        raise RuntimeError(msg)


def _raise_if_failed(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {_lib.fa_last_error().decode()}")


def _ptr(t):
    return t.data_ptr() if t is not None else None


def _round_multiple(x, m):
    return (x + m - 1) // m * m


def seqlen_k_rounded(max_seqlen_k, head_size):
    """Column extent of S_dmask.  Same rule as the reference host code for a non-sm80 device
    (fmha_api.cpp:172-179): K blocks of 128 when the padded head dim is 128, else 256."""
    blocksize_c = 128 if head_size > 64 else 256
    if max_seqlen_k <= 128:
        return 128
    if max_seqlen_k <= 256:
        return 256
    return _round_multiple(max_seqlen_k, blocksize_c)


def _check_qkv(q, k, v, cu_seqlens_q, cu_seqlens_k):
    """The TORCH_CHECKs of fmha_api.cpp:135-170.  One combined test on the hot path (a Python call per check costs
    ~1 us, 17 of them per forward and again per backward); the itemised checks only run to produce the message."""
    qs, ks, dt = q.shape, k.shape, q.dtype
    if (q.is_cuda and k.is_cuda and v.is_cuda and cu_seqlens_q.is_cuda and cu_seqlens_k.is_cuda
            and (dt is torch.float16 or dt is torch.bfloat16) and k.dtype is dt and v.dtype is dt
            and cu_seqlens_q.dtype is torch.int32 and cu_seqlens_k.dtype is torch.int32
            and len(qs) == 3 and len(ks) == 3 and v.shape == ks and ks[1:] == qs[1:]
            and q.stride(2) == 1 and k.stride(2) == 1 and v.stride(2) == 1
            and cu_seqlens_q.is_contiguous() and cu_seqlens_k.is_contiguous()
            and qs[2] % 8 == 0 and qs[2] <= 128
            and cu_seqlens_q.numel() > 1 and cu_seqlens_k.numel() == cu_seqlens_q.numel()):
        return cu_seqlens_q.numel() - 1, qs[0], ks[0], qs[1], qs[2]
    return _check_qkv_itemised(q, k, v, cu_seqlens_q, cu_seqlens_k)


def _check_qkv_itemised(q, k, v, cu_seqlens_q, cu_seqlens_k):
    _check(q.is_cuda and k.is_cuda and v.is_cuda, "q, k, v must be CUDA tensors")
    _check(q.dtype in (torch.float16, torch.bfloat16), "q must be fp16 or bf16")
    _check(k.dtype == q.dtype and v.dtype == q.dtype, "q, k, v must have the same dtype")
    _check(cu_seqlens_q.dtype == torch.int32 and cu_seqlens_k.dtype == torch.int32, "cu_seqlens must be int32")
    _check(cu_seqlens_q.is_cuda and cu_seqlens_k.is_cuda, "cu_seqlens must be CUDA tensors")
    _check(q.dim() == 3 and k.dim() == 3 and v.dim() == 3, "q, k, v must be (total, nheads, headdim)")
    _check(q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, "last dim of q, k, v must be contiguous")
    _check(cu_seqlens_q.is_contiguous() and cu_seqlens_k.is_contiguous(), "cu_seqlens must be contiguous")
    batch_size = cu_seqlens_q.numel() - 1
    _check(batch_size > 0, "batch size must be positive")
    total_q, num_heads, head_size = q.shape
    total_k = k.shape[0]
    _check(head_size % 8 == 0 and head_size <= 128, "head dim must be a multiple of 8 and at most 128")
    _check(tuple(k.shape) == (total_k, num_heads, head_size), "k has the wrong shape")
    _check(tuple(v.shape) == (total_k, num_heads, head_size), "v has the wrong shape")
    _check(cu_seqlens_k.numel() == batch_size + 1, "cu_seqlens_k has the wrong shape")
    return batch_size, total_q, total_k, num_heads, head_size


class PhiloxState:
    """The dropout stream of ONE forward call: (seed, offset) host integers and, for a call captured in a CUDA graph,
    the device counter the kernels add to the offset.  A forward hands it to its backward (as `gen`), which therefore
    regenerates exactly the same mask -- on any device, without touching the global RNG state."""
    __slots__ = ("seed", "offset", "counter")

    def __init__(self, seed, offset, counter=None):
        self.seed, self.offset, self.counter = seed, offset, counter


# Graph-capturable dropout: per device (seed, base offset, int64[1] device counter).  The CUDA generator cannot be
# queried during capture, so a captured forward takes the cached seed and lets a one-thread kernel hand out the
# counter (fa_philox_advance): every replay of the graph then draws a fresh mask, like the reference's
# PhiloxCudaState (csrc/flash_attn/src/fmha.h:39).  Filled by the first uncaptured dropout call on the device.
_graph_rng = {}


def _philox_state(p_dropout, gen, device, params, raw_stream):
    """-> PhiloxState or None.  Advances the generator like at::CUDAGeneratorImpl::philox_cuda_state
    (fmha_api.cpp:228-235).  No RNG state is touched when p_dropout == 0 (fmha_api.cpp:231)."""
    if p_dropout <= 0.0:
        return None
    if isinstance(gen, PhiloxState):
        return gen
    idx = device.index if device.index is not None else torch.cuda.current_device()
    inc = int(_lib.fa_philox_increment(ctypes.byref(params)))
    if torch.cuda.is_current_stream_capturing():
        st = _graph_rng.get(idx)
        _check(st is not None, "dropout under CUDA graph capture needs one uncaptured dropout call on this device first "
                               "(the usual warm-up iteration): it creates the device-side Philox counter")
        counter = torch.empty(1, dtype=torch.int64, device=device)
        _raise_if_failed(_lib.fa_philox_advance(st[2].data_ptr(), counter.data_ptr(), inc, raw_stream), "fa_philox_advance")
        return PhiloxState(st[0], st[1], counter)
    if gen is None:
        gen = torch.cuda.default_generators[idx]
    seed = gen.initial_seed() & 0xFFFFFFFFFFFFFFFF
    offset = gen.get_offset()
    gen.set_offset(offset + inc)
    st = _graph_rng.get(idx)
    if st is None:
        _graph_rng[idx] = [seed, offset, torch.zeros(1, dtype=torch.int64, device=device)]
    else:
        st[0], st[1] = seed, offset
    return PhiloxState(seed, offset)


def _set_philox(P, state):
    if state is not None:
        P.philox_seed, P.philox_offset = state.seed, state.offset
        if state.counter is not None:
            P.philox_offset_ptr = state.counter.data_ptr()


def _fill_common(P, q, k, v, cu_seqlens_q, cu_seqlens_k, b, h, d, total_q, total_k, max_seqlen_q, max_seqlen_k,
                 p_dropout, softmax_scale, is_causal, seqused_q=None, seqused_k=None, alibi_slopes=None):
    P.q, P.k, P.v = q.data_ptr(), k.data_ptr(), v.data_ptr()
    P.q_row_stride, P.q_head_stride, _ = q.stride()
    P.k_row_stride, P.k_head_stride, _ = k.stride()
    P.v_row_stride, P.v_head_stride, _ = v.stride()
    P.cu_seqlens_q, P.cu_seqlens_k = cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr()
    P.b, P.h, P.d, P.total_q, P.total_k = b, h, d, total_q, total_k
    P.max_seqlen_q, P.max_seqlen_k = max_seqlen_q, max_seqlen_k
    P.seqlen_q_rounded = _round_multiple(max_seqlen_q, 16)
    P.seqlen_k_rounded = seqlen_k_rounded(max_seqlen_k, d)
    P.softmax_scale, P.p_dropout = softmax_scale, p_dropout
    P.is_causal = 1 if is_causal else 0
    P.is_bf16 = 1 if q.dtype == torch.bfloat16 else 0
    P.impl = _impl
    if seqused_q is not None or seqused_k is not None or alibi_slopes is not None:
        for name, t in (("seqused_q", seqused_q), ("seqused_k", seqused_k)):
            if t is not None:
                _check(t.is_cuda and t.dtype == torch.int32 and t.is_contiguous() and t.numel() == b,
                       f"{name} must be a contiguous int32 CUDA tensor of batch_size elements")
                setattr(P, name, t.data_ptr())
        if alibi_slopes is not None:
            _check(alibi_slopes.is_cuda and alibi_slopes.dtype == torch.float32 and alibi_slopes.is_contiguous()
                   and alibi_slopes.numel() == h, "alibi_slopes must be a contiguous fp32 CUDA tensor of nheads elements")
            P.alibi_slopes = alibi_slopes.data_ptr()


class _DeviceGuard:
    """`with torch.cuda.device(dev)` costs ~15 us of Python per call; the common case (tensor on the current device) needs
    no switch at all (at::cuda::CUDAGuard, fmha_api.cpp:184, is likewise a no-op then)."""
    __slots__ = ("idx", "prev")

    def __init__(self, device):
        self.idx = device.index
        self.prev = -1

    def __enter__(self):
        cur = _current_device()
        if self.idx is not None and self.idx != cur:
            self.prev = cur
            _set_device(self.idx)
        elif self.idx is None:
            self.idx = cur
        return self

    def __exit__(self, *exc):
        if self.prev >= 0:
            _set_device(self.prev)
        return False

    def raw_stream(self):
        return _raw_stream(self.idx)   # at::cuda::getCurrentCUDAStream(), fmha_api.cpp:132


# the private fast paths of torch.cuda when this torch has them, the public API otherwise
_current_device = getattr(torch._C, "_cuda_getDevice", None) or torch.cuda.current_device
_set_device = getattr(torch._C, "_cuda_setDevice", None) or torch.cuda.set_device
if hasattr(torch._C, "_cuda_getCurrentRawStream"):
    _raw_stream = torch._C._cuda_getCurrentRawStream
else:
    def _raw_stream(idx):
        return torch.cuda.current_stream(idx).cuda_stream


# Backward scratch kept per (device, stream): [tensor, dirty_lo, dirty_hi].  The library leaves the fp32 dQ accumulator
# and its counters (the head of the workspace, fa_workspace_zero_range) all-zero when a backward completes, so a kept
# buffer is cleared once at allocation and never again (FA_FLAG_WS_ZEROED) -- no 400 MB memset per call at the headline
# size.  [dirty_lo, dirty_hi) tracks the statistics area behind the zero range, which moves when the sizes change.
_ws_cache = {}


def _workspace(P, op, device, raw_stream=None):
    nbytes = int(_lib.fa_workspace_bytes(ctypes.byref(P), op))
    if nbytes == 0:
        return None
    if op != FA_OP_BWD or raw_stream is None or torch.cuda.is_current_stream_capturing():
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)      # the library clears what it needs
        P.workspace, P.workspace_bytes = ws.data_ptr(), nbytes
        return ws
    zlen = int(_lib.fa_workspace_zero_range(ctypes.byref(P), op, None))  # the range starts at byte 0 for FA_OP_BWD
    key = (device.index, raw_stream)
    ent = _ws_cache.get(key)
    if ent is None or ent[0].numel() < nbytes:
        ent = _ws_cache[key] = [torch.zeros(nbytes, dtype=torch.uint8, device=device), nbytes, nbytes]
    ws, lo, hi = ent
    if zlen > lo:                       # an earlier, smaller call left statistics where this call accumulates
        ws[lo:min(hi, zlen)].zero_()
    ent[1], ent[2] = zlen, (max(hi, nbytes) if hi > zlen else nbytes)
    P.workspace, P.workspace_bytes = ws.data_ptr(), nbytes
    P.flags |= FA_FLAG_WS_ZEROED
    return ws


def _fwd_impl(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale,
              zero_tensors, is_causal, return_softmax, gen, blockmask=None, seqused_q=None, seqused_k=None,
              alibi_slopes=None):
    """-> ([o, lse (, s)], PhiloxState or None)"""
    b, total_q, total_k, h, d = _check_qkv(q, k, v, cu_seqlens_q, cu_seqlens_k)
    _check(p_dropout < 1.0, "p_dropout must be < 1")
    P = FaParams()
    _fill_common(P, q, k, v, cu_seqlens_q, cu_seqlens_k, b, h, d, total_q, total_k, int(max_seqlen_q),
                 int(max_seqlen_k), float(p_dropout), float(softmax_scale), is_causal, seqused_q, seqused_k, alibi_slopes)
    with _DeviceGuard(q.device) as guard:
        o = torch.empty((total_q, h, d), dtype=q.dtype, device=q.device)
        lse = torch.empty((b, h, P.seqlen_q_rounded), dtype=torch.float32, device=q.device)
        s = None
        if return_softmax:
            # zero-filled: the kernel writes only (row < seqlen_q, col < seqlen_k) entries
            s = torch.zeros((b, h, P.seqlen_q_rounded, P.seqlen_k_rounded), dtype=q.dtype, device=q.device)
            P.s_dmask = s.data_ptr()
        if zero_tensors:   # fmha_api.cpp:199-203
            o.zero_()
            lse.fill_(-math.inf)
        P.o, P.o_row_stride, P.o_head_stride = o.data_ptr(), o.stride(0), o.stride(1)
        P.softmax_lse = lse.data_ptr()
        op = FA_OP_FWD
        if blockmask is not None:
            _check(blockmask.dtype == torch.int32 and blockmask.is_cuda and blockmask.is_contiguous()
                   and blockmask.dim() == 2, "blockmask must be a contiguous 2-D int32 CUDA tensor (converted format)")
            P.blockmask = blockmask.data_ptr()
            P.blockmask_ncol, P.blockmask_nrow = blockmask.shape
            op = FA_OP_FWD_BLOCK
        stream = guard.raw_stream()
        ws = _workspace(P, op, q.device)
        state = _philox_state(p_dropout, gen, q.device, P, stream)
        _set_philox(P, state)
        # the workspace is allocated and consumed on the current stream: the caching allocator's same-stream reuse
        # rule already orders any later owner after this kernel
        fn = _lib.fa_fwd if blockmask is None else _lib.fa_fwd_block
        _raise_if_failed(fn(ctypes.byref(P), stream), "flash_attn_cuda.fwd")
    return [o, lse] + ([s] if return_softmax else []), state


def _bwd_impl(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q,
              max_seqlen_k, p_dropout, softmax_scale, zero_tensors, is_causal, gen, blockmask=None, seqused_q=None,
              seqused_k=None, alibi_slopes=None, dq_workspace=None):
    b, total_q, total_k, h, d = _check_qkv(q, k, v, cu_seqlens_q, cu_seqlens_k)
    dt, q_shape, k_shape = q.dtype, (total_q, h, d), (total_k, h, d)
    if dq_workspace is not None:      # dQ stays in the caller's workspace (FA_FLAG_KEEP_DQ_ACCUM): no dq tensor is written
        _check(dq is None and blockmask is None and not zero_tensors,
               "dq_workspace: pass dq=None (bwd_dq_finalize produces it), no blockmask, zero_tensors=False")
        dq = q                        # stands in for the shape / dtype checks below; its pointer is not handed to the library
    if not (dout.is_cuda and out.is_cuda and dq.is_cuda and dk.is_cuda and dv.is_cuda
            and dout.dtype is dt and out.dtype is dt and dq.dtype is dt and dk.dtype is dt and dv.dtype is dt
            and dout.shape == q_shape and out.shape == q_shape and dq.shape == q_shape
            and dk.shape == k_shape and dv.shape == k_shape
            and softmax_lse.dtype is torch.float32 and softmax_lse.is_contiguous()
            and dq.stride(2) == 1 and dk.stride(2) == 1 and dv.stride(2) == 1):
        # itemised only to produce the message (see _check_qkv)
        _check(dout.is_cuda and out.is_cuda and dq.is_cuda and dk.is_cuda and dv.is_cuda, "all tensors must be CUDA")
        for name, t, shape in (("dout", dout, q_shape), ("out", out, q_shape), ("dq", dq, q_shape), ("dk", dk, k_shape),
                               ("dv", dv, k_shape)):
            _check(t.dtype == q.dtype, f"{name} must have the dtype of q")
            _check(tuple(t.shape) == shape, f"{name} has the wrong shape")
        _check(softmax_lse.dtype == torch.float32 and softmax_lse.is_contiguous(), "softmax_lse must be contiguous fp32")
        _check(dq.stride(-1) == 1 and dk.stride(-1) == 1 and dv.stride(-1) == 1, "last dim of dq, dk, dv must be contiguous")
    if dout.stride(-1) != 1:
        dout = dout.contiguous()
    if out.stride(-1) != 1:
        out = out.contiguous()
    P = FaParams()
    _fill_common(P, q, k, v, cu_seqlens_q, cu_seqlens_k, b, h, d, total_q, total_k, int(max_seqlen_q),
                 int(max_seqlen_k), float(p_dropout), float(softmax_scale), is_causal, seqused_q, seqused_k, alibi_slopes)
    _check(tuple(softmax_lse.shape) == (b, h, P.seqlen_q_rounded), "softmax_lse has the wrong shape")
    with _DeviceGuard(q.device) as guard:
        softmax_d = torch.empty((b, h, P.seqlen_q_rounded), dtype=torch.float32, device=q.device)
        if zero_tensors:
            dq.zero_(), dk.zero_(), dv.zero_(), softmax_d.zero_()
        P.o, (P.o_row_stride, P.o_head_stride, _) = out.data_ptr(), out.stride()
        P.dout, (P.do_row_stride, P.do_head_stride, _) = dout.data_ptr(), dout.stride()
        P.dq, (P.dq_row_stride, P.dq_head_stride, _) = dq.data_ptr(), dq.stride()
        P.dk, (P.dk_row_stride, P.dk_head_stride, _) = dk.data_ptr(), dk.stride()
        P.dv, (P.dv_row_stride, P.dv_head_stride, _) = dv.data_ptr(), dv.stride()
        P.softmax_lse, P.softmax_d = softmax_lse.data_ptr(), softmax_d.data_ptr()
        op = FA_OP_BWD
        if blockmask is not None:
            P.blockmask = blockmask.data_ptr()
            P.blockmask_ncol, P.blockmask_nrow = blockmask.shape
            op = FA_OP_BWD_BLOCK
        stream = guard.raw_stream()
        if dq_workspace is not None:
            nbytes = int(_lib.fa_workspace_bytes(ctypes.byref(P), op))
            _check(dq_workspace.is_cuda and dq_workspace.dtype == torch.uint8 and dq_workspace.numel() >= nbytes,
                   f"dq_workspace must be a uint8 CUDA tensor of at least {nbytes} bytes (bwd_workspace_bytes), zero-filled once")
            P.workspace, P.workspace_bytes = dq_workspace.data_ptr(), dq_workspace.numel()
            P.flags |= FA_FLAG_WS_ZEROED | FA_FLAG_KEEP_DQ_ACCUM
            P.dq = None
        else:
            ws = _workspace(P, op, q.device, stream)
            if blockmask is None and P.impl != FA_IMPL_SIMT and _want_deterministic():
                P.flags |= FA_FLAG_DETERMINISTIC
        _set_philox(P, _philox_state(p_dropout, gen, q.device, P, stream))
        fn = _lib.fa_bwd if blockmask is None else _lib.fa_bwd_block
        _raise_if_failed(fn(ctypes.byref(P), stream), "flash_attn_cuda.bwd")
    return softmax_d


# ---------------------------------------------------------------------------------------------------
# the four entry points of the reference extension, same positional signatures
# ---------------------------------------------------------------------------------------------------
def fwd(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale,
        zero_tensors, is_causal, return_softmax, gen=None):
    """-> [o (total_q, h, d), softmax_lse (b, h, round16(max_seqlen_q)) fp32 (, s_dmask)].
    Signature of mha_fwd, fmha_api.cpp:112-125.  s_dmask deviates from the reference's MMA-fragment layout:
    it is the natural (b, h, sq_rounded, sk_rounded) matrix of +-P with P = exp(scale*s - lse) (true
    probabilities), sign bit set where the position was dropped."""
    return _fwd_impl(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, p_dropout,
                     softmax_scale, zero_tensors, is_causal, return_softmax, gen)[0]


def fwd_ex(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale, is_causal,
           return_softmax=False, gen=None, seqused_q=None, seqused_k=None, alibi_slopes=None):
    """`fwd` plus what the reference extension does not have: -> ([o, lse (, s_dmask)], PhiloxState or None).
    The PhiloxState names the dropout stream this call used; pass it to `bwd` as `gen` (no global RNG round trip,
    works under CUDA-graph capture).  seqused_q / seqused_k: int32 (batch,) row counts that replace the
    cu_seqlens differences (sequences may then leave gaps); alibi_slopes: fp32 (nheads,), additive bias
    -slope * |i - j| on the scaled scores."""
    return _fwd_impl(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale,
                     False, is_causal, return_softmax, gen, seqused_q=seqused_q, seqused_k=seqused_k,
                     alibi_slopes=alibi_slopes)


def bwd(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
        p_dropout, softmax_scale, zero_tensors, is_causal, gen=None, seqused_q=None, seqused_k=None,
        alibi_slopes=None, dq_workspace=None):
    """Writes dq, dk, dv in place (any row/head stride) and returns softmax_d = rowsum(dO*O).
    Call site: flash_attn/flash_attn_interface.py:31-33.  `gen` may be the PhiloxState of the forward (fwd_ex).
    dq_workspace (with dq=None): a caller-owned, once-zeroed uint8 tensor of bwd_workspace_bytes(...) bytes in which dQ keeps
    accumulating in fp32 over several calls with the same q layout (FA_FLAG_KEEP_DQ_ACCUM); bwd_dq_finalize() converts it."""
    return _bwd_impl(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q,
                     max_seqlen_k, p_dropout, softmax_scale, zero_tensors, is_causal, gen, seqused_q=seqused_q,
                     seqused_k=seqused_k, alibi_slopes=alibi_slopes, dq_workspace=dq_workspace)


def _dq_layout_params(q_like, cu_seqlens_q, max_seqlen_q, b):
    P = FaParams()
    total_q, h, d = q_like.shape
    P.b, P.h, P.d, P.total_q, P.max_seqlen_q = b, h, d, total_q, int(max_seqlen_q)
    P.cu_seqlens_q = P.cu_seqlens_k = cu_seqlens_q.data_ptr()
    P.is_bf16 = 1 if q_like.dtype == torch.bfloat16 else 0
    return P


def bwd_workspace_bytes(q_like, cu_seqlens_q, max_seqlen_q):
    """bytes of scratch a backward over this query layout needs (q_like: any (total_q, h, d) tensor)"""
    return int(_lib.fa_workspace_bytes(ctypes.byref(_dq_layout_params(q_like, cu_seqlens_q, max_seqlen_q, cu_seqlens_q.numel() - 1)), FA_OP_BWD))


def bwd_dq_finalize(dq, cu_seqlens_q, max_seqlen_q, softmax_scale, dq_workspace):
    """dq (total_q, h, d) = softmax_scale * the fp32 sums that bwd(..., dq_workspace=) calls left in dq_workspace; the
    accumulator is zero afterwards (fa_bwd_dq_finalize)."""
    _check(dq.is_cuda and dq.dim() == 3 and dq.stride(2) == 1 and dq.dtype in (torch.float16, torch.bfloat16), "bad dq")
    P = _dq_layout_params(dq, cu_seqlens_q, max_seqlen_q, cu_seqlens_q.numel() - 1)
    P.dq, (P.dq_row_stride, P.dq_head_stride, _) = dq.data_ptr(), dq.stride()
    P.softmax_scale = float(softmax_scale)
    P.workspace, P.workspace_bytes = dq_workspace.data_ptr(), dq_workspace.numel()
    with _DeviceGuard(dq.device) as guard:
        _raise_if_failed(_lib.fa_bwd_dq_finalize(ctypes.byref(P), guard.raw_stream()), "fa_bwd_dq_finalize")
    return dq


def _check_qkv_packed(qkv, blockmask):
    _check(qkv.dim() == 4 and qkv.shape[1] == 3, "qkv must be (total, 3, nheads, headdim)")
    _check(blockmask.dim() == 2, "blockmask must be 2-D (converted format, see convert_blockmask)")


def fwd_block(qkv, cu_seqlens, blockmask, p_dropout, max_s, softmax_scale, is_causal, return_softmax, gen=None):
    """-> [context, softmax_lse (, S_dmask)].  Call site: flash_attn/flash_blocksparse_attn_interface.py:45-47.
    `blockmask` is the converted (ncol, nrow) int32 mask of convert_blockmask."""
    _check_qkv_packed(qkv, blockmask)
    return _fwd_impl(qkv[:, 0], qkv[:, 1], qkv[:, 2], cu_seqlens, cu_seqlens, max_s, max_s, p_dropout,
                     softmax_scale, False, is_causal, return_softmax, gen, blockmask=blockmask)[0]


def fwd_block_ex(qkv, cu_seqlens, blockmask, p_dropout, max_s, softmax_scale, is_causal, return_softmax, gen=None):
    """fwd_block -> ([context, softmax_lse (, S_dmask)], PhiloxState or None), see fwd_ex."""
    _check_qkv_packed(qkv, blockmask)
    return _fwd_impl(qkv[:, 0], qkv[:, 1], qkv[:, 2], cu_seqlens, cu_seqlens, max_s, max_s, p_dropout,
                     softmax_scale, False, is_causal, return_softmax, gen, blockmask=blockmask)


def bwd_block(dout, qkv, out, S_dmask, softmax_lse, cu_seqlens, blockmask, p_dropout, softmax_scale, max_s,
              is_causal, gen=None):
    """-> (dqkv, dp, softmax_d).  Call site: flash_attn/flash_blocksparse_attn_interface.py:56-58 (note the
    argument order differs from fwd_block).  S_dmask is accepted for signature compatibility and unused (the
    dropout mask is regenerated from the Philox stream).  dp is a debug tensor in the reference; here None."""
    _check_qkv_packed(qkv, blockmask)
    dqkv = torch.empty_like(qkv)
    softmax_d = _bwd_impl(dout, qkv[:, 0], qkv[:, 1], qkv[:, 2], out, softmax_lse, dqkv[:, 0], dqkv[:, 1],
                          dqkv[:, 2], cu_seqlens, cu_seqlens, max_s, max_s, p_dropout, softmax_scale, False,
                          is_causal, gen, blockmask=blockmask)
    return dqkv, None, softmax_d


# ---------------------------------------------------------------------------------------------------
# varlen pack / unpack kernels used by flash_attn.bert_padding
# ---------------------------------------------------------------------------------------------------
def gather_rows(src2d, indices):
    """out[i] = src2d[indices[i]]; src2d (n, w) contiguous CUDA, indices int64 CUDA."""
    _check(src2d.is_cuda and indices.is_cuda and src2d.dim() == 2 and src2d.is_contiguous(), "bad gather input")
    _check(indices.dtype == torch.int64 and indices.is_contiguous(), "indices must be contiguous int64")
    out = torch.empty((indices.numel(), src2d.shape[1]), dtype=src2d.dtype, device=src2d.device)
    row_bytes = src2d.shape[1] * src2d.element_size()
    with _DeviceGuard(src2d.device) as guard:
        rc = _lib.fa_gather_rows(_ptr(out), _ptr(src2d), _ptr(indices), indices.numel(), row_bytes, guard.raw_stream())
    _raise_if_failed(rc, "fa_gather_rows")
    return out


def scatter_rows(src2d, indices, first_axis_dim, out=None):
    """out = zeros(first_axis_dim, w) (or the given tensor, left untouched elsewhere); out[indices[i]] = src2d[i]."""
    _check(src2d.is_cuda and indices.is_cuda and src2d.dim() == 2 and src2d.is_contiguous(), "bad scatter input")
    _check(indices.dtype == torch.int64 and indices.is_contiguous(), "indices must be contiguous int64")
    if out is None:
        out = torch.zeros((first_axis_dim, src2d.shape[1]), dtype=src2d.dtype, device=src2d.device)
    row_bytes = src2d.shape[1] * src2d.element_size()
    with _DeviceGuard(src2d.device) as guard:
        rc = _lib.fa_scatter_rows(_ptr(out), _ptr(src2d), _ptr(indices), indices.numel(), row_bytes, guard.raw_stream())
    _raise_if_failed(rc, "fa_scatter_rows")
    return out


_KIND = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


def scatter_add_rows(dst2d, src2d, indices):
    """dst2d[indices[i]] += src2d[i] in place (unique indices, negative = skip): fa_scatter_add_rows."""
    _check(dst2d.is_cuda and dst2d.dim() == 2 and dst2d.is_contiguous() and src2d.is_contiguous() and src2d.dtype == dst2d.dtype
           and src2d.shape[1] == dst2d.shape[1] and dst2d.dtype in _KIND, "bad scatter_add input")
    _check(indices.dtype == torch.int64 and indices.is_contiguous(), "indices must be contiguous int64")
    with _DeviceGuard(dst2d.device) as guard:
        rc = _lib.fa_scatter_add_rows(_ptr(dst2d), _ptr(src2d), _ptr(indices), indices.numel(), dst2d.shape[1] * dst2d.element_size(),
                                      _KIND[dst2d.dtype], guard.raw_stream())
    _raise_if_failed(rc, "fa_scatter_add_rows")
    return dst2d


def varlen_index(mask):
    """(b, s) bool CUDA mask -> indices (b*s,) int64 padded with -1, cu_seqlens (b+1,) int32, stats (2,) int32
    [longest sequence, number of valid tokens] -- all device tensors, no host synchronisation (fa_varlen_index)."""
    _check(mask.is_cuda and mask.dim() == 2, "mask must be a 2-D CUDA tensor")
    m = (mask if mask.dtype == torch.bool else mask != 0).contiguous().view(torch.uint8)
    b, s = m.shape
    indices = torch.empty(b * s, dtype=torch.int64, device=mask.device)
    cu = torch.empty(b + 1, dtype=torch.int32, device=mask.device)
    stats = torch.empty(2, dtype=torch.int32, device=mask.device)
    with _DeviceGuard(mask.device) as guard:
        rc = _lib.fa_varlen_index(_ptr(m), b, s, _ptr(indices), _ptr(cu), _ptr(stats), guard.raw_stream())
    _raise_if_failed(rc, "fa_varlen_index")
    return indices, cu, stats


def merge_partials(o_acc, lse_acc, o_blk, lse_blk, batch, seqlen, first, batch_rows=None):
    """Ring / split-KV merge (fa_merge_partials), in place: o_acc fp32 and o_blk 16 bit are (rows, h, d) tensors whose
    row of (batch entry i, position t) is i * batch_rows + t (default batch_rows = seqlen); lse_acc / lse_blk are
    (batch, h, >= seqlen) fp32 views whose last dimension is contiguous and whose (batch, h) dims are packed."""
    batch_rows = seqlen if batch_rows is None else batch_rows
    _check(o_acc.is_cuda and o_acc.dtype == torch.float32 and o_acc.is_contiguous(), "o_acc must be contiguous fp32 CUDA")
    _check(o_blk.dtype in (torch.float16, torch.bfloat16) and o_blk.is_contiguous() and o_blk.shape[1:] == o_acc.shape[1:],
           "o_blk must be contiguous fp16/bf16 with the row shape of o_acc")
    for t in (lse_acc, lse_blk):
        _check(t.dtype == torch.float32 and t.dim() == 3 and t.stride(2) == 1 and t.stride(0) == t.shape[1] * t.stride(1),
               "lse tensors must be (b, h, s) fp32 with a contiguous last dim and packed (b, h)")
    _, h, d = o_acc.shape
    need = (batch - 1) * batch_rows + seqlen
    _check(o_acc.shape[0] >= need and o_blk.shape[0] >= need, "o_acc / o_blk have too few rows")
    with _DeviceGuard(o_acc.device) as guard:
        rc = _lib.fa_merge_partials(_ptr(o_acc), _ptr(lse_acc), _ptr(o_blk), _ptr(lse_blk), batch, h, d, seqlen, batch_rows,
                                    lse_acc.stride(1), lse_blk.stride(1), 1 if o_blk.dtype == torch.bfloat16 else 0,
                                    1 if first else 0, guard.raw_stream())
    _raise_if_failed(rc, "fa_merge_partials")


def accumulate_f32(dst, src, n_batches, run_elems, dst_batch_stride, src_batch_stride):
    """dst (fp32) += src (16 bit) over n_batches runs of run_elems contiguous elements (fa_accumulate_f32)."""
    _check(dst.is_cuda and dst.dtype == torch.float32 and src.dtype in (torch.float16, torch.bfloat16), "bad dtypes")
    with _DeviceGuard(dst.device) as guard:
        rc = _lib.fa_accumulate_f32(_ptr(dst), _ptr(src), n_batches, run_elems, dst_batch_stride, src_batch_stride,
                                    1 if src.dtype == torch.bfloat16 else 0, guard.raw_stream())
    _raise_if_failed(rc, "fa_accumulate_f32")


def flag_write(flag_ptr, value, raw_stream):
    """fa_flag_write on a raw stream handle; flag_ptr: device address (possibly in a peer GPU's memory)."""
    _raise_if_failed(_lib.fa_flag_write(flag_ptr, value, raw_stream), "fa_flag_write")


def flag_wait_geq(flag_ptr, value, raw_stream, timeout_ms=0):
    _raise_if_failed(_lib.fa_flag_wait_geq(flag_ptr, value, timeout_ms, raw_stream), "fa_flag_wait_geq")


def ipc_export(tensor):
    """-> (64-byte handle of the cudaMalloc allocation containing `tensor`, byte offset of the tensor in it)"""
    buf, off = (ctypes.c_uint8 * 64)(), ctypes.c_uint64()
    with _DeviceGuard(tensor.device):
        _raise_if_failed(_lib.fa_ipc_export(tensor.data_ptr(), buf, ctypes.byref(off)), "fa_ipc_export")
    return bytes(buf), off.value


def ipc_open(device, handle):
    """map another process's allocation into `device`'s address space -> base address (int)"""
    buf, base = (ctypes.c_uint8 * 64).from_buffer_copy(handle), ctypes.c_void_p()
    with _DeviceGuard(device):
        _raise_if_failed(_lib.fa_ipc_open(buf, ctypes.byref(base)), "fa_ipc_open")
    return base.value


def enable_peer_access(device, peer_index):
    """let kernels and copies of `device` reach memory of GPU `peer_index` directly (fa_enable_peer_access)"""
    with _DeviceGuard(device):
        _raise_if_failed(_lib.fa_enable_peer_access(peer_index), "fa_enable_peer_access")


def peer_copy(dst_ptr, src_ptr, nbytes, raw_stream):
    _raise_if_failed(_lib.fa_peer_copy(dst_ptr, src_ptr, nbytes, raw_stream), "fa_peer_copy")


def rotary_qk(qkv_in, qkv_out, cos, sin, seqlen, n_rot=2, conjugate=False):
    """Rotary embedding of the first n_rot parts of a packed (rows, n_parts, h, d) tensor (fa_rotary_qk).
    qkv_out may be qkv_in (in place: only the rotated parts are touched); cos / sin are fp32 (seqlen, d/2)."""
    _check(qkv_in.is_cuda and qkv_in.dim() == 4 and qkv_in.dtype in (torch.float16, torch.bfloat16),
           "qkv must be a (rows, parts, h, d) fp16/bf16 CUDA tensor")
    _check(qkv_out.shape == qkv_in.shape and qkv_out.dtype == qkv_in.dtype and qkv_out.device == qkv_in.device,
           "output must match the input")
    rows, n_parts, h, d = qkv_in.shape
    for t in (qkv_in, qkv_out):
        _check(t.stride(3) == 1 and t.stride(2) == d and t.stride(1) == h * d, "(parts, h, d) must be contiguous within a row")
    for t in (cos, sin):
        _check(t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and tuple(t.shape) == (seqlen, d // 2),
               "cos / sin must be contiguous fp32 (seqlen, d/2)")
    with _DeviceGuard(qkv_in.device) as guard:
        rc = _lib.fa_rotary_qk(_ptr(qkv_in), _ptr(qkv_out), _ptr(cos), _ptr(sin), rows, seqlen, n_parts, n_rot, h, d,
                               qkv_in.stride(0), qkv_out.stride(0), 1 if conjugate else 0,
                               1 if qkv_in.dtype == torch.bfloat16 else 0, guard.raw_stream())
    _raise_if_failed(rc, "fa_rotary_qk")
    return qkv_out


def kernel_timing(enable):
    """Benchmarks: bracket the main fwd / bwd kernels with CUDA events on the launching stream."""
    _lib.fa_timing_enable(1 if enable else 0)


def kernel_timing_collect():
    """-> {'fwd_ms', 'fwd_launches', 'bwd_ms', 'bwd_launches'} summed since the last collect (synchronises)."""
    f, b = ctypes.c_double(), ctypes.c_double()
    nf, nb = ctypes.c_uint64(), ctypes.c_uint64()
    _raise_if_failed(_lib.fa_timing_collect(ctypes.byref(f), ctypes.byref(nf), ctypes.byref(b), ctypes.byref(nb)),
                     "fa_timing_collect")
    return {"fwd_ms": f.value, "fwd_launches": nf.value, "bwd_ms": b.value, "bwd_launches": nb.value}


def dropout_keep8_host(seed, offset, bh, qrow, kgroup, p_dropout):
    """CPU evaluation of the dropout stream (tests only): keep flags of keys 8*kgroup..8*kgroup+7."""
    buf = (ctypes.c_uint8 * 8)()
    _lib.fa_dropout_keep8_host(seed, offset, bh, qrow, kgroup, p_dropout, buf)
    return [bool(x) for x in buf]
