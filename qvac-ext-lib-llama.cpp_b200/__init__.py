"""ggml-b200 — Python view of the C-ABI in include/ggml_b200.h (ctypes), for tests and bench.py.

The product is the pair of shared libraries next to this file:
  libb200mm.so     sm_100a CUDA kernels + C-ABI           (csrc/*.cu)
  libggml-b200.so  the ggml backend over that C-ABI        (csrc/ggml_backend_b200.cpp)
PyTorch is used only as a device-memory allocator and stream provider. Nothing in this package computes on the CPU and
nothing here imports oracle/: if the CUDA library is missing or there is no sm_100 device, calls raise.

The directory name contains '-' and '.', so import it through `b200_pkg.load()` at the repository root.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libb200mm.so"
BACKEND_PATH = HERE / "libggml-b200.so"

Q4_0, Q8_0, Q4_K, Q6_K = 2, 8, 12, 14
Q5_K = 13
TYPE_NAMES = {Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q5_K: "q5_K", Q6_K: "q6_K"}
BLOCK = {Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210)}
PATH_AUTO, PATH_MMVQ, PATH_MMQ, PATH_SIMPLE = 0, 1, 2, 3

EXPORTS = [
    "b200_device_count", "b200_device_info", "b200_device_memory", "b200_ctx_create", "b200_ctx_destroy", "b200_ctx_device",
    "b200_ctx_stream", "b200_ctx_synchronize", "b200_last_error", "b200_ctx_launch_count", "b200_ctx_set_force_path", "b200_ctx_set_exact",
    "b200_malloc", "b200_free", "b200_host_malloc", "b200_host_free", "b200_memcpy_h2d", "b200_memcpy_d2h", "b200_memcpy_d2d",
    "b200_memcpy_peer", "b200_memset", "b200_memcpy_h2d_async", "b200_memcpy_d2h_async", "b200_memcpy_d2d_async",
    "b200_memcpy2d_async", "b200_enable_peer_access", "b200_memcpy_peer_async",
    "b200_graph_capture_begin", "b200_graph_capture_end", "b200_graph_launch", "b200_graph_destroy", "b200_graph_kernel_count",
    "b200_event_create", "b200_event_destroy", "b200_event_record", "b200_event_wait", "b200_event_synchronize",
    "b200_event_elapsed_ms", "b200_quantize_q8_0", "b200_quantize_q8_1", "b200_quantize_q8_K", "b200_quantize_units",
    "b200_mul_mat_supported", "b200_mul_mat", "b200_mul_mat_shared_src1", "b200_mul_mat_host", "b200_mul_mat_id_supported",
    "b200_mul_mat_id", "b200_mm_ids_sort", "b200_ctx_set_row_split", "b200_ctx_set_row_split_ll", "b200_ctx_set_row_split_deferred", "b200_ctx_row_split_wait", "b200_ctx_set_weight_cache", "b200_ctx_set_mmq_f16", "b200_ctx_set_prefetch", "b200_ctx_trace", "b200_ctx_trace_read", "b200_weight_cache_invalidate", "b200_weight_cache_clear", "b200_weight_cache_bytes", "b200_weight_cache_set_budget",
    "b200_mul_mat_fused_supported", "b200_mul_mat_fused", "b200_mul_mat_id_fused_supported", "b200_mul_mat_id_fused",
    "b200_glu", "b200_add", "b200_mul", "b200_add_id",
    "b200_mul_mat_noalloc", "b200_mul_mat_noalloc_max_units",
    "b200_rms_norm_supported", "b200_rms_norm", "b200_rms_norm_mul_supported", "b200_rms_norm_mul", "b200_attn_prep_supported", "b200_attn_prep", "b200_attn_decode_supported", "b200_attn_decode", "b200_qkv_rope", "b200_flash_attn_ext_decode",
    "b200_mul_mat_shared_src1_norm", "b200_mul_mat_fused_norm", "b200_get_rows_supported", "b200_get_rows", "b200_rope_supported", "b200_rope",
    "b200_set_rows_supported", "b200_set_rows", "b200_cpy_supported", "b200_cpy", "b200_flash_attn_ext_supported", "b200_flash_attn_ext",
]


def row_size(t: int, k: int) -> int:
    e, b = BLOCK[t]
    if k % e:
        raise ValueError(f"K={k} is not a multiple of the {TYPE_NAMES[t]} block size {e}")
    return k // e * b


class MulMatDesc(C.Structure):
    _fields_ = [("type", C.c_int32),
                ("ne00", C.c_int64), ("ne01", C.c_int64), ("ne02", C.c_int64), ("ne03", C.c_int64),
                ("nb01", C.c_size_t), ("nb02", C.c_size_t), ("nb03", C.c_size_t),
                ("ne11", C.c_int64), ("ne12", C.c_int64), ("ne13", C.c_int64),
                ("nb11", C.c_size_t), ("nb12", C.c_size_t), ("nb13", C.c_size_t),
                ("nb1", C.c_size_t), ("nb2", C.c_size_t), ("nb3", C.c_size_t)]

    @classmethod
    def contiguous(cls, t, M, K, N, ne02=1, ne03=1, ne12=None, ne13=None):
        ne12 = ne02 if ne12 is None else ne12
        ne13 = ne03 if ne13 is None else ne13
        rb = row_size(t, K)
        return cls(t, K, M, ne02, ne03, rb, rb * M, rb * M * ne02, N, ne12, ne13, 4 * K, 4 * K * N, 4 * K * N * ne12,
                   4 * M, 4 * M * N, 4 * M * N * ne12)


class MulMatIdDesc(C.Structure):
    _fields_ = [("type", C.c_int32),
                ("ne00", C.c_int64), ("ne01", C.c_int64), ("n_as", C.c_int64),
                ("nb01", C.c_size_t), ("nb02", C.c_size_t),
                ("ne11", C.c_int64), ("n_tok", C.c_int64),
                ("nb11", C.c_size_t), ("nb12", C.c_size_t),
                ("n_used", C.c_int64), ("ids_nb1", C.c_size_t),
                ("nb1", C.c_size_t), ("nb2", C.c_size_t)]

    @classmethod
    def contiguous(cls, t, M, K, n_as, n_used, ne11, n_tok):
        rb = row_size(t, K)
        return cls(t, K, M, n_as, rb, rb * M, ne11, n_tok, 4 * K, 4 * K * ne11, n_used, 4 * n_used, 4 * M, 4 * M * n_used)


GLU_NONE, GLU_REGLU, GLU_GEGLU, GLU_SWIGLU, GLU_SWIGLU_OAI, GLU_GEGLU_ERF, GLU_GEGLU_QUICK = -1, 0, 1, 2, 3, 4, 5


class Fusion(C.Structure):
    """b200_fusion (include/ggml_b200.h): the K7 epilogue of a one-column MUL_MAT / one-token MUL_MAT_ID"""
    _fields_ = [("glu_op", C.c_int32), ("alpha", C.c_float), ("limit", C.c_float),
                ("up_bias", C.c_void_p), ("gate_bias", C.c_void_p),
                ("bias_nb1", C.c_size_t), ("bias_nb2", C.c_size_t), ("bias_nb3", C.c_size_t)]


class Tensor(C.Structure):
    """b200_tensor (include/ggml_b200.h): a ggml tensor's data pointer, type, ne (elements) and nb (bytes)"""
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4)]

    @classmethod
    def of(cls, t, gtype=None):
        """a torch tensor of up to 4 dims in torch order [ne3, ne2, ne1, ne0] (any strides) as a ggml tensor"""
        import torch
        types = {torch.float32: 0, torch.float16: 1, torch.int32: 26, torch.int64: 27}
        gt = types[t.dtype] if gtype is None else gtype
        t4 = t
        while t4.dim() < 4:
            t4 = t4.unsqueeze(0)
        es = t.element_size()
        return cls(t.data_ptr(), gt, (C.c_int64 * 4)(*reversed(t4.shape)), (C.c_size_t * 4)(*[st * es for st in reversed(t4.stride())]))

    @classmethod
    def quantized(cls, data, gtype, K, M):
        """a [K, M] weight matrix of ggml type `gtype` stored as bytes in `data` (torch uint8), rows contiguous"""
        rb = row_size(gtype, K)
        return cls(data.data_ptr(), gtype, (C.c_int64 * 4)(K, M, 1, 1), (C.c_size_t * 4)(BLOCK[gtype][1], rb, rb * M, rb * M))


class Norm(C.Structure):
    """b200_norm: the RMS_NORM (+ MUL) a b200_mul_mat_*_norm call applies to its src1"""
    _fields_ = [("w", C.c_void_p), ("eps", C.c_float), ("y", C.c_void_p), ("y_nb1", C.c_size_t)]


class RopeParams(C.Structure):
    """b200_rope_params: op_params[1, 2, 4..10] of a GGML_OP_ROPE node"""
    _fields_ = [("n_dims", C.c_int32), ("mode", C.c_int32), ("n_ctx_orig", C.c_int32), ("freq_base", C.c_float), ("freq_scale", C.c_float),
                ("ext_factor", C.c_float), ("attn_factor", C.c_float), ("beta_fast", C.c_float), ("beta_slow", C.c_float)]


class RowSplit(C.Structure):
    _fields_ = [("n_ranks", C.c_int32), ("rank", C.c_int32), ("dst_delta", C.c_int64 * 8), ("sync_words", C.c_void_p * 8)]


class RowSplitLL(C.Structure):
    _fields_ = [("stage", C.c_void_p * 8), ("dst_base", C.c_void_p), ("region_bytes", C.c_uint64), ("tensor_split", C.c_double * 8),
                ("rounding", C.c_int32)]


ROW_SPLIT_ROUNDING = 128   # rows: the tensor-core kernel's tile height (the reference rounds to its mmq_y, ggml-cuda.cu:773-784)


def row_split_bounds(nrows: int, n_ranks: int, rank: int, tensor_split=None, rounding: int = ROW_SPLIT_ROUNDING):
    """[row_low, row_high) of `rank`'s slice of a row-split weight tensor. Same rule as the reference's get_row_split
    (ggml-cuda.cu:786-799): tensor_split holds the CUMULATIVE start fraction of every device (llama.cpp normalises
    --tensor-split into that form, ggml-cuda.cu:941-967), boundaries are rounded down to `rounding`, the last device takes the rest.
    The slices of ranks 0..n-1 are disjoint, ordered and cover [0, nrows); a slice may be empty."""
    if tensor_split is None:
        tensor_split = [r / n_ranks for r in range(n_ranks)]
    if len(tensor_split) != n_ranks or tensor_split[0] != 0 or any(b < a for a, b in zip(tensor_split, tensor_split[1:])):
        raise ValueError("tensor_split must hold n_ranks non-decreasing cumulative fractions starting at 0")
    lo = 0 if rank == 0 else int(nrows * tensor_split[rank])
    lo -= lo % rounding
    if rank == n_ranks - 1:
        hi = nrows
    else:
        hi = int(nrows * tensor_split[rank + 1])
        hi -= hi % rounding
    return lo, hi


class B200Error(RuntimeError):
    pass


_lib = None


def lib() -> C.CDLL:
    """the C-ABI library; raises if it has not been built (no fallback of any kind)"""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise B200Error(f"{LIB_PATH} is missing: run __graft_entry__.build() (make -C {HERE / 'csrc'})")
        L = C.CDLL(str(LIB_PATH))
        vp, i64, sz, ci = C.c_void_p, C.c_int64, C.c_size_t, C.c_int
        L.b200_device_count.restype = ci
        L.b200_last_error.restype = C.c_char_p
        L.b200_ctx_create.argtypes = [ci, vp, C.POINTER(vp)]
        L.b200_ctx_destroy.argtypes = [vp]
        L.b200_ctx_destroy.restype = None
        L.b200_ctx_stream.argtypes = [vp]
        L.b200_ctx_stream.restype = vp
        L.b200_ctx_synchronize.argtypes = [vp]
        L.b200_ctx_launch_count.argtypes = [vp]
        L.b200_ctx_launch_count.restype = C.c_uint64
        L.b200_ctx_set_force_path.argtypes = [vp, ci]
        L.b200_ctx_set_force_path.restype = None
        L.b200_ctx_set_exact.argtypes = [vp, ci]
        L.b200_ctx_set_exact.restype = None
        L.b200_device_info.argtypes = [ci, C.c_char_p, sz, C.c_char_p, sz, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)]
        L.b200_host_malloc.argtypes = [sz, C.POINTER(vp)]
        L.b200_host_free.argtypes = [vp]
        for n in ("b200_quantize_q8_0", "b200_quantize_q8_1", "b200_quantize_q8_K"):
            getattr(L, n).argtypes = [vp, vp, i64, i64, sz, vp]
        L.b200_quantize_units.argtypes = [vp, ci, vp, i64, i64, sz, vp, vp]
        tp = C.POINTER(Tensor)
        L.b200_rms_norm_supported.argtypes = [tp, tp]
        L.b200_rms_norm.argtypes = [vp, tp, tp, C.c_float]
        L.b200_rms_norm_mul_supported.argtypes = [tp, tp, tp]
        L.b200_rms_norm_mul.argtypes = [vp, tp, tp, tp, C.c_float]
        rpp = C.POINTER(RopeParams)
        L.b200_attn_prep_supported.argtypes = [tp, tp, tp, tp, tp, tp, rpp, tp, tp, tp, tp, tp]
        L.b200_attn_prep.argtypes = [vp, tp, tp, tp, tp, tp, tp, rpp, tp, tp, tp, tp, tp]
        L.b200_qkv_rope.argtypes = [vp, C.POINTER(MulMatDesc), C.POINTER(vp), vp, C.POINTER(Norm), C.POINTER(vp), tp, tp, tp, tp, tp, tp, rpp, tp, tp, tp, tp, tp]
        L.b200_attn_decode_supported.argtypes = [tp, tp, tp, tp, tp, tp, rpp, tp, tp, tp, tp, tp, tp, tp, tp, tp, tp, tp]
        L.b200_attn_decode.argtypes = [vp, tp, tp, tp, tp, tp, tp, rpp, tp, tp, tp, tp, tp, tp, tp, tp, tp, tp, tp, C.c_float, C.c_float, C.c_float]
        L.b200_get_rows_supported.argtypes = [tp, tp, tp]
        L.b200_get_rows.argtypes = [vp, tp, tp, tp]
        L.b200_rope_supported.argtypes = [tp, tp, tp, tp, C.POINTER(RopeParams)]
        L.b200_rope.argtypes = [vp, tp, tp, tp, tp, C.POINTER(RopeParams)]
        L.b200_set_rows_supported.argtypes = [tp, tp, tp]
        L.b200_set_rows.argtypes = [vp, tp, tp, tp]
        L.b200_cpy_supported.argtypes = [tp, tp]
        L.b200_cpy.argtypes = [vp, tp, tp]
        L.b200_flash_attn_ext_supported.argtypes = [tp, tp, tp, tp, tp, tp]
        L.b200_flash_attn_ext.argtypes = [vp, tp, tp, tp, tp, tp, tp, C.c_float, C.c_float, C.c_float]
        L.b200_flash_attn_ext_decode.argtypes = [vp, tp, tp, tp, tp, tp, tp, C.c_float, C.c_float, C.c_float, tp]
        L.b200_mul_mat_supported.argtypes = [C.POINTER(MulMatDesc)]
        L.b200_mul_mat.argtypes = [vp, C.POINTER(MulMatDesc), vp, vp, vp]
        L.b200_mul_mat_noalloc.argtypes = [vp, C.POINTER(MulMatDesc), vp, vp, vp]
        L.b200_mul_mat_noalloc_max_units.restype = sz
        L.b200_mul_mat_shared_src1.argtypes = [vp, ci, C.POINTER(MulMatDesc), C.POINTER(vp), vp, C.POINTER(vp)]
        L.b200_mul_mat_shared_src1_norm.argtypes = [vp, ci, C.POINTER(MulMatDesc), C.POINTER(vp), vp, C.POINTER(Norm), C.POINTER(vp)]
        L.b200_mul_mat_fused_norm.argtypes = [vp, C.POINTER(MulMatDesc), vp, vp, vp, C.POINTER(Norm), C.POINTER(Fusion), vp]
        L.b200_mul_mat_host.argtypes = [vp, C.POINTER(MulMatDesc), vp, ci, vp, vp]
        L.b200_mul_mat_id_supported.argtypes = [C.POINTER(MulMatIdDesc)]
        L.b200_mul_mat_id.argtypes = [vp, C.POINTER(MulMatIdDesc), vp, vp, vp, vp]
        L.b200_mm_ids_sort.argtypes = [vp, vp, i64, i64, sz, i64, vp, vp]
        L.b200_mul_mat_fused_supported.argtypes = [vp, C.POINTER(MulMatDesc)]
        L.b200_mul_mat_fused.argtypes = [vp, C.POINTER(MulMatDesc), vp, vp, vp, C.POINTER(Fusion), vp]
        L.b200_mul_mat_id_fused_supported.argtypes = [vp, C.POINTER(MulMatIdDesc)]
        L.b200_mul_mat_id_fused.argtypes = [vp, C.POINTER(MulMatIdDesc), vp, vp, vp, vp, C.POINTER(Fusion), vp]
        L.b200_glu.argtypes = [vp, ci, C.c_float, C.c_float, vp, sz, vp, sz, vp, sz, i64, i64]
        L.b200_add.argtypes = [vp, C.POINTER(i64), vp, C.POINTER(sz), vp, C.POINTER(i64), C.POINTER(sz), vp, C.POINTER(sz)]
        L.b200_mul.argtypes = L.b200_add.argtypes
        L.b200_add_id.argtypes = [vp, vp, sz, sz, vp, sz, i64, vp, sz, vp, sz, sz, i64, i64, i64]
        for n in ("b200_memcpy_h2d_async", "b200_memcpy_d2h_async", "b200_memcpy_d2d_async"):
            getattr(L, n).argtypes = [vp, vp, vp, sz]
        L.b200_ctx_set_row_split.argtypes = [vp, C.POINTER(RowSplit)]
        L.b200_ctx_set_weight_cache.argtypes = [vp, ci]
        L.b200_ctx_set_weight_cache.restype = None
        L.b200_ctx_set_row_split_ll.argtypes = [vp, vp]
        L.b200_ctx_set_row_split_ll.restype = ci
        L.b200_ctx_set_mmq_f16.argtypes = [vp, ci]
        L.b200_ctx_set_mmq_f16.restype = None
        L.b200_ctx_set_row_split_deferred.argtypes = [vp, ci]
        L.b200_ctx_row_split_wait.argtypes = [vp]
        L.b200_ctx_set_prefetch.argtypes = [vp, ci]
        L.b200_ctx_set_prefetch.restype = None
        L.b200_ctx_trace.argtypes = [vp, ci]
        L.b200_ctx_trace_read.argtypes = [vp, vp, C.c_size_t]
        L.b200_weight_cache_invalidate.argtypes = [ci, vp, sz]
        L.b200_weight_cache_invalidate.restype = None
        L.b200_weight_cache_clear.argtypes = [ci]
        L.b200_weight_cache_clear.restype = None
        L.b200_weight_cache_bytes.argtypes = []
        L.b200_weight_cache_bytes.restype = sz
        _lib = L
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise B200Error(f"{what} failed (status {rc}): {lib().b200_last_error().decode(errors='replace')}")


class Context:
    """one device + one stream (a ggml_backend_t). By default launches on torch's current stream so torch events see it."""

    def __init__(self, device: int = 0, use_torch_stream: bool = True):
        import torch
        if not torch.cuda.is_available():
            raise B200Error("no CUDA device: ggml-b200 has no CPU fallback")
        self.torch = torch
        self.device = device
        torch.cuda.set_device(device)
        # torch's default stream has handle 0, which the C-ABI reads as "create a stream": name the legacy default stream
        # explicitly (cudaStreamLegacy == 1) so that torch's copies and .cpu() reads stay ordered with our launches
        stream = (torch.cuda.current_stream(device).cuda_stream or 1) if use_torch_stream else None
        h = C.c_void_p()
        check(lib().b200_ctx_create(device, C.c_void_p(stream) if stream else None, C.byref(h)), "b200_ctx_create")
        self.h = h

    def close(self):
        if self.h:
            lib().b200_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib().b200_ctx_synchronize(self.h), "synchronize")

    @property
    def launches(self) -> int:
        return int(lib().b200_ctx_launch_count(self.h))

    def force_path(self, p: int):
        lib().b200_ctx_set_force_path(self.h, p)

    def exact(self, enable: bool):
        """every MUL_MAT(_ID) on the exact-order kernel: bit-identical to ggml-cpu's generic build (validation only)"""
        lib().b200_ctx_set_exact(self.h, 1 if enable else 0)

    def row_split(self, rank: int, dst_ptrs, sync_ptrs):
        """fused all-gather mode: dst_ptrs[r] / sync_ptrs[r] = base address of the symmetric dst region / sync words on rank r"""
        n = len(dst_ptrs)
        if n <= 1:
            check(lib().b200_ctx_set_row_split(self.h, None), "b200_ctx_set_row_split")
            return
        rs = RowSplit()
        rs.n_ranks, rs.rank = n, rank
        for r in range(n):
            rs.dst_delta[r] = int(dst_ptrs[r]) - int(dst_ptrs[rank])
            rs.sync_words[r] = int(sync_ptrs[r])
        check(lib().b200_ctx_set_row_split(self.h, C.byref(rs)), "b200_ctx_set_row_split")

    def row_split_ll(self, stage_ptrs, dst_base: int, region_bytes: int, tensor_split=None, rounding: int = ROW_SPLIT_ROUNDING):
        """flagged-payload exchange for the decode mat-vec kernels (b200_ctx_set_row_split_ll): stage_ptrs[r] = rank r's staging area
        (4 * region_bytes bytes, zeroed, peer-mapped), dst_base / region_bytes = the local symmetric dst region. None switches it off."""
        if stage_ptrs is None:
            check(lib().b200_ctx_set_row_split_ll(self.h, None), "b200_ctx_set_row_split_ll")
            return
        n = len(stage_ptrs)
        ll = RowSplitLL()
        for r in range(n):
            ll.stage[r] = int(stage_ptrs[r])
            ll.tensor_split[r] = (r / n) if tensor_split is None else float(tensor_split[r])
        ll.dst_base, ll.region_bytes, ll.rounding = int(dst_base), int(region_bytes), int(rounding)
        check(lib().b200_ctx_set_row_split_ll(self.h, C.byref(ll)), "b200_ctx_set_row_split_ll")

    def row_split_deferred(self, enable: bool = True):
        """decode kernels publish their slice and let the next launch (or row_split_wait) wait for the peers"""
        check(lib().b200_ctx_set_row_split_deferred(self.h, 1 if enable else 0), "b200_ctx_set_row_split_deferred")

    def row_split_wait(self):
        check(lib().b200_ctx_row_split_wait(self.h), "b200_ctx_row_split_wait")

    def weight_cache(self, enable: bool):
        """keep int8 expansions of weight matrices across calls; the caller must invalidate ranges it overwrites or frees"""
        lib().b200_ctx_set_weight_cache(self.h, 1 if enable else 0)

    def mmq_f16(self, enable: bool):
        """prefill arithmetic: True = kind::f16 GEMM on scale-carrying fp16 operands (default), False = exact per-block kind::i8"""
        lib().b200_ctx_set_mmq_f16(self.h, 1 if enable else 0)

    def prefetch(self, kb_per_cta=32):
        """cross-kernel weight prefetch for decode (b200_ctx_set_prefetch); 0 switches it off"""
        lib().b200_ctx_set_prefetch(self.h, int(kb_per_cta))

    def trace(self, n_slots: int):
        """debug timeline of the mat-vec launches (b200_ctx_trace); 0 = off"""
        check(lib().b200_ctx_trace(self.h, int(n_slots)), "b200_ctx_trace")
        self._trace_slots = int(n_slots)

    def trace_read(self, sm_count: int = 148):
        """-> uint64 array [n_slots, sm_count, 8]"""
        import numpy as np
        a = np.zeros((self._trace_slots, sm_count, 8), dtype=np.uint64)
        check(lib().b200_ctx_trace_read(self.h, a.ctypes.data, a.nbytes), "b200_ctx_trace_read")
        return a

    def weight_cache_clear(self):
        lib().b200_weight_cache_clear(self.device)

    # ---- iface.set_tensor_async / get_tensor_async on the context's stream
    def h2d_async(self, dst_dev_ptr: int, src_host_ptr: int, nbytes: int):
        check(lib().b200_memcpy_h2d_async(self.h, dst_dev_ptr, src_host_ptr, nbytes), "b200_memcpy_h2d_async")

    def d2h_async(self, dst_host_ptr: int, src_dev_ptr: int, nbytes: int):
        check(lib().b200_memcpy_d2h_async(self.h, dst_host_ptr, src_dev_ptr, nbytes), "b200_memcpy_d2h_async")

    # ---- K1
    def quantize(self, name: str, x):
        """x: float32 cuda tensor [rows, k] (last dim contiguous) -> uint8 tensor [rows, row_bytes] in the reference block layout"""
        torch = self.torch
        sz = {"q8_0": (32, 34), "q8_1": (32, 36), "q8_K": (256, 292)}[name]
        rows, k = x.shape
        y = torch.empty((rows, k // sz[0] * sz[1]), dtype=torch.uint8, device=x.device)
        f = getattr(lib(), f"b200_quantize_{name}")
        check(f(self.h, x.data_ptr(), k, rows, x.stride(0) * 4, y.data_ptr()), f"b200_quantize_{name}")
        return y

    def quantize_units(self, kind: int, x):
        torch = self.torch
        rows, k = x.shape
        u = k // 256
        main = torch.empty((rows, u, 272), dtype=torch.uint8, device=x.device)
        aux = torch.empty((rows, u, 16), dtype=torch.uint8, device=x.device)
        check(lib().b200_quantize_units(self.h, kind, x.data_ptr(), k, rows, x.stride(0) * 4, main.data_ptr(), aux.data_ptr()),
              "b200_quantize_units")
        return main, aux

    # ---- K3/K4
    def mul_mat(self, t: int, W, X, out=None, desc: MulMatDesc | None = None):
        """W: uint8 cuda tensor holding ggml blocks [M, row_bytes] (or [ne03, ne02, M, row_bytes]);
        X: float32 [N, K] (or [ne13, ne12, N, K]); returns float32 [N, M] (or [ne13, ne12, N, M])"""
        torch = self.torch
        K = X.shape[-1]
        if desc is None:
            if X.dim() == 2:
                M = W.numel() // row_size(t, K)
                desc = MulMatDesc.contiguous(t, M, K, X.shape[0])
                oshape = (X.shape[0], M)
            else:
                ne03, ne02, M = W.shape[0], W.shape[1], W.shape[2]
                ne13, ne12, N = X.shape[0], X.shape[1], X.shape[2]
                desc = MulMatDesc.contiguous(t, M, K, N, ne02, ne03, ne12, ne13)
                oshape = (ne13, ne12, N, M)
        else:
            oshape = (desc.ne13, desc.ne12, desc.ne11, desc.ne01)
        if out is None:
            out = torch.empty(oshape, dtype=torch.float32, device=X.device)
        check(lib().b200_mul_mat(self.h, C.byref(desc), W.data_ptr(), X.data_ptr(), out.data_ptr()), "b200_mul_mat")
        return out

    def mul_mat_shared(self, t: int, Ws, X, outs=None, dst_strides=None, norm_w=None, norm_eps=None):
        """several weight matrices against one X in one launch; dst_strides[i] = bytes between dst columns of output i
        (default: contiguous [N, M_i]). norm_eps: X is un-normalized, the call computes RMS_NORM(X) (* norm_w) first - inside the
        mat-vec kernel where it can (b200_mul_mat_shared_src1_norm)"""
        torch = self.torch
        N, K = X.shape
        n = len(Ws)
        descs = (MulMatDesc * n)()
        Ms = []
        for i, W in enumerate(Ws):
            M = W.numel() // row_size(t, K)
            Ms.append(M)
            descs[i] = MulMatDesc.contiguous(t, M, K, N)
            if dst_strides is not None:
                descs[i].nb1 = dst_strides[i]
                descs[i].nb2 = descs[i].nb3 = dst_strides[i] * N
        if outs is None:
            outs = [torch.empty((N, M), dtype=torch.float32, device=X.device) for M in Ms]
        wp = (C.c_void_p * n)(*[W.data_ptr() for W in Ws])
        op = (C.c_void_p * n)(*[o.data_ptr() for o in outs])
        if norm_eps is not None:
            y = torch.empty_like(X)
            nm = Norm(None if norm_w is None else norm_w.data_ptr(), norm_eps, y.data_ptr(), K * 4)
            check(lib().b200_mul_mat_shared_src1_norm(self.h, n, descs, wp, X.data_ptr(), C.byref(nm), op), "b200_mul_mat_shared_src1_norm")
            return outs
        check(lib().b200_mul_mat_shared_src1(self.h, n, descs, wp, X.data_ptr(), op), "b200_mul_mat_shared_src1")
        return outs

    def mul_mat_host(self, desc: MulMatDesc, W_dev_ptr: int, x_host_ptr: int, y_host_ptr: int, w_on_device: bool = True):
        check(lib().b200_mul_mat_host(self.h, C.byref(desc), W_dev_ptr, 1 if w_on_device else 0, x_host_ptr, y_host_ptr), "b200_mul_mat_host")

    # ---- K5/K6
    def mul_mat_id(self, t: int, W, B, ids, out=None):
        """W uint8 [n_as, M, row_bytes]; B float32 [n_tok, ne11, K]; ids int32 [n_tok, n_used] -> float32 [n_tok, n_used, M]"""
        torch = self.torch
        n_tok, ne11, K = B.shape
        n_as, M = W.shape[0], W.shape[1]
        n_used = ids.shape[1]
        desc = MulMatIdDesc.contiguous(t, M, K, n_as, n_used, ne11, n_tok)
        if out is None:
            out = torch.empty((n_tok, n_used, M), dtype=torch.float32, device=B.device)
        check(lib().b200_mul_mat_id(self.h, C.byref(desc), W.data_ptr(), B.data_ptr(), ids.data_ptr(), out.data_ptr()), "b200_mul_mat_id")
        return out

    # ---- K7: fused epilogues and the stand-alone element-wise nodes
    def mul_mat_fused(self, t: int, W_up, W_gate, X, up_bias=None, gate_bias=None, glu_op: int = GLU_NONE, alpha=1.702, limit=7.0, out=None,
                      norm_w=None, norm_eps=None):
        """one column: X float32 [1, K]; W_* uint8 [M, row_bytes]; biases float32 [M] -> float32 [1, M] in ONE launch"""
        torch = self.torch
        K = X.shape[-1]
        M = W_up.numel() // row_size(t, K)
        desc = MulMatDesc.contiguous(t, M, K, 1)
        f = Fusion(glu_op, alpha, limit, None if up_bias is None else up_bias.data_ptr(), None if gate_bias is None else gate_bias.data_ptr(),
                   0, 4 * M, 4 * M)
        if out is None:
            out = torch.empty((1, M), dtype=torch.float32, device=X.device)
        if norm_eps is not None:
            y = torch.empty_like(X)
            nm = Norm(None if norm_w is None else norm_w.data_ptr(), norm_eps, y.data_ptr(), K * 4)
            check(lib().b200_mul_mat_fused_norm(self.h, C.byref(desc), W_up.data_ptr(), None if W_gate is None else W_gate.data_ptr(), X.data_ptr(),
                                                C.byref(nm), C.byref(f), out.data_ptr()), "b200_mul_mat_fused_norm")
            return out
        check(lib().b200_mul_mat_fused(self.h, C.byref(desc), W_up.data_ptr(), None if W_gate is None else W_gate.data_ptr(), X.data_ptr(),
                                       C.byref(f), out.data_ptr()), "b200_mul_mat_fused")
        return out

    def mul_mat_id_fused(self, t: int, W_up, W_gate, B, ids, up_bias=None, gate_bias=None, glu_op: int = GLU_NONE, alpha=1.702, limit=7.0, out=None):
        """one token: W_* uint8 [n_as, M, row_bytes]; B float32 [1, ne11, K]; ids int32 [1, n_used]; biases float32 [n_as, M]"""
        torch = self.torch
        n_tok, ne11, K = B.shape
        n_as, M = W_up.shape[0], W_up.shape[1]
        n_used = ids.shape[1]
        desc = MulMatIdDesc.contiguous(t, M, K, n_as, n_used, ne11, n_tok)
        f = Fusion(glu_op, alpha, limit, None if up_bias is None else up_bias.data_ptr(), None if gate_bias is None else gate_bias.data_ptr(),
                   4 * M, 0, 0)
        if out is None:
            out = torch.empty((n_tok, n_used, M), dtype=torch.float32, device=B.device)
        check(lib().b200_mul_mat_id_fused(self.h, C.byref(desc), W_up.data_ptr(), None if W_gate is None else W_gate.data_ptr(), B.data_ptr(),
                                          ids.data_ptr(), C.byref(f), out.data_ptr()), "b200_mul_mat_id_fused")
        return out

    def glu(self, op: int, gate, up, alpha=1.702, limit=7.0):
        """GGML_OP_GLU, split form: gate / up float32 [rows, nc] (row strides honoured) -> [rows, nc]"""
        out = self.torch.empty((gate.shape[0], gate.shape[1]), dtype=self.torch.float32, device=gate.device)
        check(lib().b200_glu(self.h, op, alpha, limit, gate.data_ptr(), gate.stride(0) * 4, up.data_ptr(), up.stride(0) * 4, out.data_ptr(),
                             out.stride(0) * 4, gate.shape[1], gate.shape[0]), "b200_glu")
        return out

    def mul(self, a, b):
        """GGML_OP_MUL, same broadcasting as add"""
        return self.add(a, b, fn="b200_mul")

    def add(self, a, b, fn="b200_add"):
        """GGML_OP_ADD with ggml broadcasting: a, b float32 tensors of up to 4 dims in torch order [ne3, ne2, ne1, ne0]"""
        a4 = a.reshape((1,) * (4 - a.dim()) + tuple(a.shape))
        b4 = b.reshape((1,) * (4 - b.dim()) + tuple(b.shape))
        out = self.torch.empty_like(a4)
        i64x4, szx4 = C.c_int64 * 4, C.c_size_t * 4
        ne = lambda x: i64x4(*reversed(x.shape))
        nb = lambda x: szx4(*[s * 4 for s in reversed(x.stride())])
        check(getattr(lib(), fn)(self.h, ne(a4), a4.data_ptr(), nb(a4), b4.data_ptr(), ne(b4), nb(b4), out.data_ptr(), nb(out)), fn)
        return out.reshape(a.shape)

    # ---- f3: the ops between the matmuls (ops.cu). Tensors are torch tensors in torch order [ne3, ne2, ne1, ne0], any strides.
    def rms_norm(self, x, eps=1e-6, out=None):
        """GGML_OP_RMS_NORM over the last torch dim"""
        out = self.torch.empty(x.shape, dtype=self.torch.float32, device=x.device) if out is None else out
        check(lib().b200_rms_norm(self.h, Tensor.of(x), Tensor.of(out), eps), "b200_rms_norm")
        return out

    def rms_norm_mul(self, x, w, eps=1e-6):
        """RMS_NORM followed by MUL with the per-channel weight w (broadcast over rows), one launch"""
        out = self.torch.empty(x.shape, dtype=self.torch.float32, device=x.device)
        check(lib().b200_rms_norm_mul(self.h, Tensor.of(x), Tensor.of(w), Tensor.of(out), eps), "b200_rms_norm_mul")
        return out

    def attn_prep(self, q, k, pos, k_cache, k_idx, v, v_cache, v_idx, n_dims, mode=0, freq_factors=None, **kw):
        """ROPE(q), ROPE(k) + SET_ROWS(k_cache), SET_ROWS(v_cache) in one launch; returns (rope(q), rope(k)). q [1, n_tok, n_head, D],
        k [1, n_tok, n_head_kv, D], v [n_tok, n_head_kv * D], caches [n_ctx, n_head_kv * D] float16 / float32"""
        qo = self.torch.empty(q.shape, dtype=self.torch.float32, device=q.device)
        ko = self.torch.empty(k.shape, dtype=self.torch.float32, device=k.device)
        rp = RopeParams(n_dims, mode, kw.get("n_ctx_orig", 0), kw.get("freq_base", 10000.0), kw.get("freq_scale", 1.0), kw.get("ext_factor", 0.0),
                        kw.get("attn_factor", 1.0), kw.get("beta_fast", 32.0), kw.get("beta_slow", 1.0))
        check(lib().b200_attn_prep(self.h, Tensor.of(q), Tensor.of(qo), Tensor.of(k), Tensor.of(ko), Tensor.of(pos),
                                   None if freq_factors is None else Tensor.of(freq_factors), rp, Tensor.of(k_idx), Tensor.of(k_cache), Tensor.of(v),
                                   Tensor.of(v_idx), Tensor.of(v_cache)), "b200_attn_prep")
        return qo, ko

    def qkv_rope(self, t, Ws, x, pos, k_cache, k_idx, v_cache, v_idx, n_head, n_head_kv, n_dims, mode=0, norm_w=None, norm_eps=None, freq_factors=None, **kw):
        """[RMS_NORM + MUL] + q/k/v projections + ROPE(q), ROPE(k) + the two KV-cache rows: x float32 [n_tok, K] -> (rope(q) [1, n_tok, n_head, D],
        rope(k) [1, n_tok, n_head_kv, D]); one launch for one token (b200_qkv_rope)"""
        torch = self.torch
        N, K = x.shape
        descs = (MulMatDesc * 3)()
        Ms = [W.numel() // row_size(t, K) for W in Ws]
        D = Ms[0] // n_head
        for i in range(3):
            descs[i] = MulMatDesc.contiguous(t, Ms[i], K, N)
        raw = [torch.empty((N, M), dtype=torch.float32, device=x.device) for M in Ms]
        wp = (C.c_void_p * 3)(*[W.data_ptr() for W in Ws])
        op = (C.c_void_p * 3)(*[o.data_ptr() for o in raw])
        q, k, v = raw[0].view(1, N, n_head, D), raw[1].view(1, N, n_head_kv, D), raw[2]
        qo, ko = torch.empty_like(q), torch.empty_like(k)
        nm = None
        if norm_eps is not None:
            y = torch.empty_like(x)
            nm = C.byref(Norm(None if norm_w is None else norm_w.data_ptr(), norm_eps, y.data_ptr(), K * 4))
        rp = RopeParams(n_dims, mode, kw.get("n_ctx_orig", 0), kw.get("freq_base", 10000.0), kw.get("freq_scale", 1.0), kw.get("ext_factor", 0.0),
                        kw.get("attn_factor", 1.0), kw.get("beta_fast", 32.0), kw.get("beta_slow", 1.0))
        check(lib().b200_qkv_rope(self.h, descs, wp, x.data_ptr(), nm, op, Tensor.of(q), Tensor.of(qo), Tensor.of(k), Tensor.of(ko), Tensor.of(pos),
                                  None if freq_factors is None else Tensor.of(freq_factors), rp, Tensor.of(k_idx), Tensor.of(k_cache), Tensor.of(v),
                                  Tensor.of(v_idx), Tensor.of(v_cache)), "b200_qkv_rope")
        return qo, ko

    def attn_decode(self, q, k, pos, k_cache, k_idx, v, v_cache, v_idx, n_kv, mask, n_dims, mode=0, scale=1.0, freq_factors=None, sinks=None, **kw):
        """one token: attn_prep + flash_attn_ext in ONE launch. q [1, 1, n_head, D], k [1, 1, n_head_kv, D], v [1, n_head_kv * D]; caches
        [n_ctx, n_head_kv * D] float16; attention over cache rows [0, n_kv). Returns (attention [1, 1, n_head, D], rope(q), rope(k))"""
        torch = self.torch
        nh, D = q.shape[2], q.shape[3]
        nkv = k.shape[2]
        qo = torch.empty(q.shape, dtype=torch.float32, device=q.device)
        ko = torch.empty(k.shape, dtype=torch.float32, device=k.device)
        out = torch.empty((1, 1, nh, D), dtype=torch.float32, device=q.device)
        rp = RopeParams(n_dims, mode, kw.get("n_ctx_orig", 0), kw.get("freq_base", 10000.0), kw.get("freq_scale", 1.0), kw.get("ext_factor", 0.0),
                        kw.get("attn_factor", 1.0), kw.get("beta_fast", 32.0), kw.get("beta_slow", 1.0))
        fa_q = qo.permute(0, 2, 1, 3)
        fa_k = k_cache[:n_kv].view(1, n_kv, nkv, D).permute(0, 2, 1, 3)
        fa_v = v_cache[:n_kv].view(1, n_kv, nkv, D).permute(0, 2, 1, 3)
        check(lib().b200_attn_decode(self.h, Tensor.of(q), Tensor.of(qo), Tensor.of(k), Tensor.of(ko), Tensor.of(pos),
                                     None if freq_factors is None else Tensor.of(freq_factors), rp, Tensor.of(k_idx), Tensor.of(k_cache), Tensor.of(v),
                                     Tensor.of(v_idx), Tensor.of(v_cache), Tensor.of(fa_q), Tensor.of(fa_k), Tensor.of(fa_v),
                                     None if mask is None else Tensor.of(mask), None if sinks is None else Tensor.of(sinks), Tensor.of(out), scale,
                                     kw.get("max_bias", 0.0), kw.get("logit_softcap", 0.0)), "b200_attn_decode")
        return out, qo, ko

    def get_rows(self, src0, ids, out=None):
        """GGML_OP_GET_ROWS: src0 a torch tensor [.., n_rows, K] (float32 / float16) or a Tensor (quantized); ids int32 [.., n_ids]"""
        t0 = src0 if isinstance(src0, Tensor) else Tensor.of(src0)
        ti = Tensor.of(ids)
        if out is None:
            out = self.torch.empty((ti.ne[2], ti.ne[1], ti.ne[0], t0.ne[0]), dtype=self.torch.float32, device=ids.device)
        check(lib().b200_get_rows(self.h, t0, ti, Tensor.of(out)), "b200_get_rows")
        return out

    def rope(self, x, pos, n_dims, mode=0, freq_factors=None, n_ctx_orig=0, freq_base=10000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0,
             beta_fast=32.0, beta_slow=1.0, out=None):
        """GGML_OP_ROPE: x float32 [n_batch, n_tokens, n_head, n_embd_head]; pos int32 [n_tokens]"""
        out = self.torch.empty(x.shape, dtype=self.torch.float32, device=x.device) if out is None else out
        rp = RopeParams(n_dims, mode, n_ctx_orig, freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow)
        check(lib().b200_rope(self.h, Tensor.of(x), Tensor.of(pos), None if freq_factors is None else Tensor.of(freq_factors), Tensor.of(out), rp), "b200_rope")
        return out

    def set_rows(self, dst, src, idx):
        """GGML_OP_SET_ROWS: dst[.., idx[i], :] = src[.., i, :]; src float32, dst float32 / float16, idx int64 / int32"""
        check(lib().b200_set_rows(self.h, Tensor.of(src), Tensor.of(idx), Tensor.of(dst)), "b200_set_rows")
        return dst

    def cpy(self, src, dst):
        """GGML_OP_CPY / CONT / DUP between float32 / float16 tensors of equal element count, any strides"""
        check(lib().b200_cpy(self.h, Tensor.of(src), Tensor.of(dst)), "b200_cpy")
        return dst

    def flash_attn_ext(self, q, k, v, mask=None, sinks=None, scale=1.0, max_bias=0.0, logit_softcap=0.0, written_idx=None):
        """GGML_OP_FLASH_ATTN_EXT: q float32 [n_batch, n_head, n_q, D]; k, v float16 [n_batch, n_head_kv, n_kv, D]; mask float16
        [1 | n_batch, 1 | n_head, >= n_q, n_kv]; returns float32 [n_batch, n_q, n_head, D]"""
        nb_, nh, nq, D = q.shape
        out = self.torch.empty((nb_, nq, nh, D), dtype=self.torch.float32, device=q.device)
        if written_idx is not None:   # one token behind its own KV-cache write: b200_flash_attn_ext_decode
            check(lib().b200_flash_attn_ext_decode(self.h, Tensor.of(q), Tensor.of(k), Tensor.of(v), None if mask is None else Tensor.of(mask),
                                                   None if sinks is None else Tensor.of(sinks), Tensor.of(out), scale, max_bias, logit_softcap,
                                                   Tensor.of(written_idx)), "b200_flash_attn_ext_decode")
            return out
        check(lib().b200_flash_attn_ext(self.h, Tensor.of(q), Tensor.of(k), Tensor.of(v), None if mask is None else Tensor.of(mask),
                                        None if sinks is None else Tensor.of(sinks), Tensor.of(out), scale, max_bias, logit_softcap), "b200_flash_attn_ext")
        return out

    def add_id(self, a, bias, ids):
        """GGML_OP_ADD_ID: a float32 [n_tok, n_used, M]; bias [n_as, M]; ids int32 [n_tok, n_used]"""
        out = self.torch.empty_like(a)
        n_tok, n_used, M = a.shape
        check(lib().b200_add_id(self.h, a.data_ptr(), a.stride(1) * 4, a.stride(0) * 4, bias.data_ptr(), bias.stride(0) * 4, bias.shape[0],
                                ids.data_ptr(), ids.stride(0) * 4, out.data_ptr(), out.stride(1) * 4, out.stride(0) * 4, M, n_used, n_tok), "b200_add_id")
        return out

    def mm_ids_sort(self, ids, n_as: int):
        torch = self.torch
        n_tok, n_used = ids.shape
        sp = torch.empty(n_tok * n_used, dtype=torch.int32, device=ids.device)
        eb = torch.empty(n_as + 1, dtype=torch.int32, device=ids.device)
        check(lib().b200_mm_ids_sort(self.h, ids.data_ptr(), n_used, n_tok, 4 * n_used, n_as, sp.data_ptr(), eb.data_ptr()), "b200_mm_ids_sort")
        return sp, eb


def device_count() -> int:
    return int(lib().b200_device_count())
