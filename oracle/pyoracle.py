"""oracle/pyoracle.py — TEST INFRASTRUCTURE ONLY (not product code).

ctypes bindings for
  * oracle/_ref/liboracle_c.so   — our plain-C restatement (oracle/oracle.c), class `Oracle`
  * oracle/_ref/libggml-base.so  — the REFERENCE's own quantizers, class `RefQuant`
    (quantize_row_q8_0_ref / q8_1_ref / q8_K_ref, ggml_quantize_chunk; /root/reference/ggml/src/ggml-quants.c)
  * oracle/_ref/libref_harness.so — one-op graphs through the reference's public backend API, class `RefHarness`

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg may import this module.
Nothing here reads /root/reference at run time: the .so files were built from it by oracle/Makefile
and travel with the repository snapshot.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"

Q4_0, Q8_0, Q4_K, Q6_K = 2, 8, 12, 14
Q5_K = 13   # enum ggml_type values (ggml/include/ggml.h:378-395)
GGML_TYPE_Q8_1, GGML_TYPE_Q8_K = 9, 15
TYPE_NAMES = {Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q5_K: "q5_K", Q6_K: "q6_K"}
BLOCK = {Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210)}   # (elements, bytes)


def row_size(t: int, k: int) -> int:
    e, b = BLOCK[t]
    assert k % e == 0
    return k // e * b


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """our C restatement"""

    def __init__(self):
        p = REF_DIR / "liboracle_c.so"
        if not p.exists():
            raise FileNotFoundError(f"{p} missing — run `make -C oracle oracle` (or __graft_entry__.build())")
        L = self.lib = C.CDLL(str(p))
        L.orc_fp32_to_fp16.restype = C.c_uint16
        L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_fp16_to_fp32.restype = C.c_float
        L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
        for n in ("q8_0", "q8_1", "q8_K", "q4_0"):
            f = getattr(L, f"orc_quantize_row_{n}")
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        for n in ("q4_0", "q8_0", "q4_K", "q5_K", "q6_K"):
            f = getattr(L, f"orc_dequantize_row_{n}")
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        for n in ("q4_0_q8_0", "q8_0_q8_0", "q4_K_q8_K", "q5_K_q8_K", "q6_K_q8_K"):
            f = getattr(L, f"orc_vec_dot_{n}")
            f.restype = C.c_float
            f.argtypes = [C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_mul_mat.restype = None
        L.orc_mul_mat.argtypes = [C.c_int, C.c_void_p] + [C.c_int64] * 4 + [C.c_size_t] * 3 + \
                                 [C.c_void_p] + [C.c_int64] * 3 + [C.c_size_t] * 3 + [C.c_void_p]
        L.orc_mul_mat_id.restype = None
        L.orc_mul_mat_id.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_size_t, C.c_size_t,
                                     C.c_void_p, C.c_int64, C.c_int64, C.c_size_t, C.c_size_t,
                                     C.c_void_p, C.c_int64, C.c_size_t, C.c_void_p]
        L.orc_nmse.restype = C.c_double
        L.orc_nmse.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_glu.restype = None
        L.orc_glu.argtypes = [C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_add_row.restype = None
        L.orc_add_row.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.orc_add_id.restype = None
        L.orc_add_id.argtypes = [C.c_void_p] * 4 + [C.c_int64] * 4

    # -- quantizers: x float32 [rows, k] -> uint8 [rows, row_bytes]
    def quantize(self, name: str, x: np.ndarray) -> np.ndarray:
        sz = {"q8_0": (32, 34), "q8_1": (32, 36), "q8_K": (256, 292), "q4_0": (32, 18)}[name]
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.reshape(-1, x.shape[-1]).shape
        out = np.zeros((rows, k // sz[0] * sz[1]), dtype=np.uint8)
        f = getattr(self.lib, f"orc_quantize_row_{name}")
        xr = x.reshape(rows, k)
        for r in range(rows):
            f(_ptr(xr[r]), _ptr(out[r]), k)
        return out

    def dequantize(self, t: int, w: np.ndarray, k: int) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, row_size(t, k))
        out = np.zeros((w.shape[0], k), dtype=np.float32)
        f = getattr(self.lib, f"orc_dequantize_row_{TYPE_NAMES[t]}")
        for r in range(w.shape[0]):
            f(_ptr(w[r]), _ptr(out[r]), k)
        return out

    def vec_dot(self, t: int, k: int, w_row: np.ndarray, a_row: np.ndarray) -> float:
        n = {Q4_0: "q4_0_q8_0", Q8_0: "q8_0_q8_0", Q4_K: "q4_K_q8_K", Q5_K: "q5_K_q8_K", Q6_K: "q6_K_q8_K"}[t]
        return float(getattr(self.lib, f"orc_vec_dot_{n}")(k, _ptr(np.ascontiguousarray(w_row)), _ptr(np.ascontiguousarray(a_row))))

    def mul_mat(self, t: int, W: np.ndarray, X: np.ndarray, ne02: int = 1, ne03: int = 1) -> np.ndarray:
        """W uint8 [ne03, ne02, M, row_bytes] (or [M,row_bytes]); X float32 [ne13, ne12, N, K] (or [N,K]) -> [ne13,ne12,N,M]"""
        X = np.ascontiguousarray(X, dtype=np.float32)
        W = np.ascontiguousarray(W, dtype=np.uint8)
        K = X.shape[-1]
        X4 = X.reshape((1, 1) + X.shape)[..., :, :].reshape((-1,) + X.shape[-2:]) if X.ndim == 2 else X
        if X.ndim == 2:
            ne12 = ne13 = 1
            N = X.shape[0]
        else:
            assert X.ndim == 4
            ne13, ne12, N, _ = X.shape
        rb = row_size(t, K)
        M = W.size // (rb * ne02 * ne03)
        out = np.zeros((ne13, ne12, N, M), dtype=np.float32)
        self.lib.orc_mul_mat(t, _ptr(W), M, K, ne02, ne03, rb, rb * M, rb * M * ne02,
                             _ptr(X), N, ne12, ne13, 4 * K, 4 * K * N, 4 * K * N * ne12, _ptr(out))
        return out if X.ndim == 4 else out[0, 0]

    def mul_mat_id(self, t: int, W: np.ndarray, B: np.ndarray, ids: np.ndarray) -> np.ndarray:
        """W uint8 [n_as, M, row_bytes]; B float32 [n_tok, ne11, K]; ids int32 [n_tok, n_used] -> [n_tok, n_used, M]"""
        B = np.ascontiguousarray(B, dtype=np.float32)
        W = np.ascontiguousarray(W, dtype=np.uint8)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n_tok, ne11, K = B.shape
        n_as, M, rb = W.shape
        assert rb == row_size(t, K)
        n_used = ids.shape[1]
        out = np.zeros((n_tok, n_used, M), dtype=np.float32)
        self.lib.orc_mul_mat_id(t, _ptr(W), M, K, n_as, rb, rb * M, _ptr(B), ne11, n_tok, 4 * K, 4 * K * ne11,
                                _ptr(ids), n_used, 4 * n_used, _ptr(out))
        return out

    # -- K7: the element-wise nodes around the MUL_MATs
    def glu(self, op: int, gate: np.ndarray, up: np.ndarray, alpha: float = 1.702, limit: float = 7.0) -> np.ndarray:
        gate = np.ascontiguousarray(gate, dtype=np.float32)
        up = np.ascontiguousarray(up, dtype=np.float32)
        assert gate.shape == up.shape
        out = np.zeros_like(gate)
        self.lib.orc_glu(op, alpha, limit, _ptr(gate), _ptr(up), _ptr(out), gate.size)
        return out

    def add_row(self, a: np.ndarray, bias: np.ndarray) -> np.ndarray:
        """a float32 [..., M] + bias [M]"""
        a = np.ascontiguousarray(a, dtype=np.float32)
        bias = np.ascontiguousarray(bias, dtype=np.float32)
        out = np.zeros_like(a)
        self.lib.orc_add_row(_ptr(a), _ptr(bias), _ptr(out), a.shape[-1], a.size // a.shape[-1])
        return out

    def add_id(self, a: np.ndarray, bias: np.ndarray, ids: np.ndarray) -> np.ndarray:
        """a float32 [n_tok, n_used, M]; bias [n_as, M]; ids int32 [n_tok, n_used]"""
        a = np.ascontiguousarray(a, dtype=np.float32)
        bias = np.ascontiguousarray(bias, dtype=np.float32)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        out = np.zeros_like(a)
        self.lib.orc_add_id(_ptr(a), _ptr(bias), _ptr(ids), _ptr(out), a.shape[2], a.shape[1], a.shape[0], bias.shape[0])
        return out

    def mul_mat_glu(self, t: int, W_up, W_gate, X, up_bias=None, gate_bias=None, glu_op: int = -1, alpha=1.702, limit=7.0) -> np.ndarray:
        """the fused patterns of test_mul_mat_vec_fusion as a composition of the node oracles: X [N, K] -> [N, M]"""
        up = self.mul_mat(t, W_up, X)
        if up_bias is not None:
            up = self.add_row(up, up_bias)
        if glu_op < 0:
            return up
        gate = self.mul_mat(t, W_gate, X)
        if gate_bias is not None:
            gate = self.add_row(gate, gate_bias)
        return self.glu(glu_op, gate, up, alpha, limit)

    def mul_mat_id_glu(self, t: int, W_up, W_gate, B, ids, up_bias=None, gate_bias=None, glu_op: int = -1, alpha=1.702, limit=7.0) -> np.ndarray:
        up = self.mul_mat_id(t, W_up, B, ids)
        if up_bias is not None:
            up = self.add_id(up, up_bias, ids)
        if glu_op < 0:
            return up
        gate = self.mul_mat_id(t, W_gate, B, ids)
        if gate_bias is not None:
            gate = self.add_id(gate, gate_bias, ids)
        return self.glu(glu_op, gate, up, alpha, limit)

    def nmse(self, ref: np.ndarray, got: np.ndarray) -> float:
        a = np.ascontiguousarray(ref, dtype=np.float32).ravel()
        b = np.ascontiguousarray(got, dtype=np.float32).ravel()
        assert a.size == b.size
        return float(self.lib.orc_nmse(_ptr(a), _ptr(b), a.size))


class RefQuant:
    """the reference's own quantizers from oracle/_ref/libggml-base.so"""

    def __init__(self):
        p = REF_DIR / "libggml-base.so"
        if not p.exists():
            raise FileNotFoundError(f"{p} missing — run `make -C oracle ref` where /root/reference exists")
        L = self.lib = C.CDLL(str(p))
        for n in ("q8_0", "q8_1", "q8_K", "q4_0", "q4_K"):
            f = getattr(L, f"quantize_row_{n}_ref")
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.ggml_quantize_chunk.restype = C.c_size_t
        L.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
        for n in ("q4_0", "q8_0", "q4_K", "q5_K", "q6_K"):
            f = getattr(L, f"dequantize_row_{n}")
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]

    def quantize_ref(self, name: str, x: np.ndarray) -> np.ndarray:
        sz = {"q8_0": (32, 34), "q8_1": (32, 36), "q8_K": (256, 292), "q4_0": (32, 18), "q4_K": (256, 144)}[name]
        x = np.ascontiguousarray(x, dtype=np.float32)
        xr = x.reshape(-1, x.shape[-1])
        rows, k = xr.shape
        out = np.zeros((rows, k // sz[0] * sz[1]), dtype=np.uint8)
        f = getattr(self.lib, f"quantize_row_{name}_ref")
        for r in range(rows):
            f(_ptr(xr[r]), _ptr(out[r]), k)
        return out

    def quantize_weights(self, t: int, w: np.ndarray) -> np.ndarray:
        """ggml_quantize_chunk(type, f32, dst, 0, nrows, K, imatrix=NULL) (ggml.c:7442) -> uint8 [rows, row_bytes]"""
        w = np.ascontiguousarray(w, dtype=np.float32)
        wr = w.reshape(-1, w.shape[-1])
        rows, k = wr.shape
        out = np.zeros((rows, row_size(t, k)), dtype=np.uint8)
        n = self.lib.ggml_quantize_chunk(t, _ptr(wr), _ptr(out), 0, rows, k, None)
        assert n == out.size
        return out

    def dequantize(self, t: int, w: np.ndarray, k: int) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, row_size(t, k))
        out = np.zeros((w.shape[0], k), dtype=np.float32)
        f = getattr(self.lib, f"dequantize_row_{TYPE_NAMES[t]}")
        for r in range(w.shape[0]):
            f(_ptr(w[r]), _ptr(out[r]), k)
        return out


class RefHarness:
    """one-op graphs through the reference's public ggml-backend API (oracle/ref_harness.cpp)"""

    _inst = None

    def __init__(self):
        for n in ("libggml-base.so", "libggml.so", "libref_harness.so"):
            if not (REF_DIR / n).exists():
                raise FileNotFoundError(f"{REF_DIR / n} missing — run `make -C oracle ref` where /root/reference exists")
        # RTLD_LOCAL: libref_harness.so finds both through DT_NEEDED + rpath; nothing of ggml's goes into the global scope
        L = self.lib = C.CDLL(str(REF_DIR / "libref_harness.so"))
        L.refh_load_backend.restype = C.c_int
        L.refh_load_backend.argtypes = [C.c_char_p]
        L.refh_reg_name.restype = C.c_char_p
        L.refh_reg_name.argtypes = [C.c_int]
        L.refh_reg_dev_count.restype = C.c_int
        L.refh_reg_dev_count.argtypes = [C.c_int]
        L.refh_dev_description.restype = C.c_char_p
        L.refh_dev_description.argtypes = [C.c_int, C.c_int]
        L.refh_mul_mat.restype = C.c_int
        L.refh_mul_mat.argtypes = [C.c_int] * 4 + [C.c_int64] * 7 + [C.c_void_p] * 3 + [C.c_double, C.POINTER(C.c_double)]
        L.refh_mul_mat_split.restype = C.c_int
        L.refh_mul_mat_split.argtypes = [C.c_int] * 3 + [C.c_int64] * 3 + [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_int]
        L.refh_mul_mat_id.restype = C.c_int
        L.refh_mul_mat_id.argtypes = [C.c_int] * 4 + [C.c_int64] * 6 + [C.c_void_p] * 4 + [C.c_double, C.POINTER(C.c_double)]
        L.refh_mul_mat_stack.restype = C.c_int
        L.refh_proc_u64.restype = C.c_uint64
        L.refh_proc_u64.argtypes = [C.c_int, C.c_char_p]
        L.refh_mul_mat_glu.restype = C.c_int
        L.refh_mul_mat_glu.argtypes = [C.c_int] * 4 + [C.c_int64] * 3 + [C.c_void_p] * 5 + [C.c_int, C.c_float, C.c_float, C.c_void_p]
        L.refh_mul_mat_id_glu.restype = C.c_int
        L.refh_mul_mat_id_glu.argtypes = [C.c_int] * 4 + [C.c_int64] * 6 + [C.c_void_p] * 6 + [C.c_int, C.c_float, C.c_float, C.c_void_p]
        L.refh_mul_mat_stack.argtypes = [C.c_int] * 5 + [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                                        C.c_double, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.refh_stack_e2e.restype = C.c_int
        L.refh_stack_e2e.argtypes = [C.c_int] * 5 + [C.c_void_p] * 3 + [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                     C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self._loaded = {}

    @classmethod
    def get(cls) -> "RefHarness":
        if cls._inst is None:
            cls._inst = cls()
        return cls._inst

    def load(self, path) -> int:
        path = str(path)
        if path not in self._loaded:
            h = self.lib.refh_load_backend(path.encode())
            if h < 0:
                raise RuntimeError(f"ggml_backend_load({path}) failed")
            self._loaded[path] = h
        return self._loaded[path]

    def cpu(self, variant: str = "x64") -> int:
        """variant: 'x64' (generic scalar code, strict oracle), 'haswell' (AVX2) or 'icelake' (AVX512-VNNI)"""
        return self.load(REF_DIR / f"libggml-cpu-{variant}.so")

    def best_cpu_variant(self) -> str:
        flags = ""
        try:
            with open("/proc/cpuinfo") as f:
                for line in f:
                    if line.startswith("flags"):
                        flags = line
                        break
        except OSError:
            pass
        fl = set(flags.split())
        if {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512vbmi", "avx512_vnni"} <= fl:
            return "icelake"
        if {"avx2", "fma", "f16c", "bmi2"} <= fl:
            return "haswell"
        return "x64"

    def mul_mat(self, reg: int, t: int, W: np.ndarray, X: np.ndarray, ne02=1, ne03=1, n_threads=0, dev=0,
                seconds=0.0):
        """W uint8 [ne03*ne02*M, row_bytes]; X float32 [ne13, ne12, N, K] or [N, K]. returns (out, us_per_run|None)"""
        X = np.ascontiguousarray(X, dtype=np.float32)
        W = np.ascontiguousarray(W, dtype=np.uint8)
        K = X.shape[-1]
        if X.ndim == 2:
            ne13 = ne12 = 1
            N = X.shape[0]
        else:
            ne13, ne12, N, _ = X.shape
        M = W.size // (row_size(t, K) * ne02 * ne03)
        out = np.zeros((ne13, ne12, N, M), dtype=np.float32)
        us = C.c_double(0)
        rc = self.lib.refh_mul_mat(reg, dev, n_threads, t, M, K, N, ne02, ne03, ne12, ne13,
                                   _ptr(W), _ptr(X), _ptr(out), seconds, C.byref(us))
        if rc:
            raise RuntimeError(f"refh_mul_mat rc={rc} (2 = backend does not support the op)")
        out = out if X.ndim == 4 else out[0, 0]
        return out, (us.value if seconds > 0 else None)

    def stack_e2e(self, reg: int, t: int, shapes, x_of, x_K, N, W_all, X_all, warmup=3, steps=10, n_threads=0, dev=0, want_out=True):
        """the MUL_MAT stack of one step as ONE cgraph through ggml_backend_tensor_set_async / graph_compute_async / tensor_get_async /
        synchronize (refh_stack_e2e). shapes: [(M, K)]; x_of[i]: which activation tensor matrix i reads; x_K: K of every activation
        tensor. returns (us_compute, us_e2e, out float32 [sum(M) * N] | None)"""
        Ms = np.ascontiguousarray([m for m, _ in shapes], dtype=np.int64)
        Ks = np.ascontiguousarray([k for _, k in shapes], dtype=np.int64)
        xo = np.ascontiguousarray(x_of, dtype=np.int32)
        xk = np.ascontiguousarray(x_K, dtype=np.int64)
        W_all = np.ascontiguousarray(W_all, dtype=np.uint8)
        X_all = np.ascontiguousarray(X_all, dtype=np.float32)
        assert X_all.size == int(xk.sum()) * N and len(xo) == len(Ms)
        out = np.zeros(int(Ms.sum()) * N, dtype=np.float32) if want_out else None
        uc, ue = C.c_double(0), C.c_double(0)
        rc = self.lib.refh_stack_e2e(reg, dev, n_threads, t, len(Ms), _ptr(Ms), _ptr(Ks), _ptr(xo), len(xk), _ptr(xk), N, _ptr(W_all), _ptr(X_all),
                                     warmup, steps, None if out is None else _ptr(out), C.byref(uc), C.byref(ue))
        if rc:
            raise RuntimeError(f"refh_stack_e2e rc={rc} (2 = op not supported, 3 = allocation failed, 4 = graph_compute failed)")
        return uc.value, ue.value, out

    def mul_mat_split(self, reg: int, t: int, W: np.ndarray, X: np.ndarray, tensor_split=None, dev=0, repeats=1):
        """MUL_MAT with W in the backend's split buffer type (ggml_backend_split_buffer_type): W uint8 [M, row_bytes], X float32
        [N, K]; tensor_split = proportions per device (None: the backend's default). returns (out [N, M], weight bytes read back)"""
        X = np.ascontiguousarray(X, dtype=np.float32)
        W = np.ascontiguousarray(W, dtype=np.uint8)
        N, K = X.shape
        M = W.size // row_size(t, K)
        out = np.zeros((N, M), dtype=np.float32)
        back = np.zeros_like(W)
        ts = None if tensor_split is None else np.ascontiguousarray(tensor_split, dtype=np.float32)
        rc = self.lib.refh_mul_mat_split(reg, dev, t, M, K, N, _ptr(W), _ptr(X), _ptr(out), None if ts is None else _ptr(ts),
                                         0 if ts is None else len(ts), _ptr(back), repeats)
        if rc:
            raise RuntimeError(f"refh_mul_mat_split rc={rc} (2 = op not supported, 5 = no split buffer type)")
        return out, back

    def mul_mat_id(self, reg: int, t: int, W: np.ndarray, B: np.ndarray, ids: np.ndarray, n_threads=0, dev=0,
                   seconds=0.0):
        """W uint8 [n_as, M, row_bytes]; B float32 [n_tok, ne11, K]; ids int32 [n_tok, n_used]"""
        B = np.ascontiguousarray(B, dtype=np.float32)
        W = np.ascontiguousarray(W, dtype=np.uint8)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n_tok, ne11, K = B.shape
        n_as, M, _ = W.shape
        n_used = ids.shape[1]
        out = np.zeros((n_tok, n_used, M), dtype=np.float32)
        us = C.c_double(0)
        rc = self.lib.refh_mul_mat_id(reg, dev, n_threads, t, M, K, n_as, n_used, ne11, n_tok,
                                      _ptr(W), _ptr(B), _ptr(ids), _ptr(out), seconds, C.byref(us))
        if rc:
            raise RuntimeError(f"refh_mul_mat_id rc={rc}")
        return out, (us.value if seconds > 0 else None)

    def proc_u64(self, reg: int, name: str):
        """value of a `uint64_t (*)(void)` the registry exposes via get_proc_address; None if it has no such entry"""
        v = int(self.lib.refh_proc_u64(reg, name.encode()))
        return None if v == 2 ** 64 - 1 else v

    def mul_mat_glu(self, reg: int, t: int, W_up, W_gate, X, up_bias=None, gate_bias=None, glu_op: int = -1, alpha=1.702, limit=7.0,
                    n_threads=0, dev=0) -> np.ndarray:
        """{MUL_MAT [, ADD], MUL_MAT [, ADD], GLU} (glu_op >= 0) or {MUL_MAT, ADD} as one graph on backend `reg`: X [N, K] -> [N, M]"""
        X = np.ascontiguousarray(X, dtype=np.float32)
        N, K = X.shape
        W_up = np.ascontiguousarray(W_up, dtype=np.uint8)
        M = W_up.size // row_size(t, K)
        opt = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)
        W_gate, up_bias, gate_bias = opt(W_gate, np.uint8), opt(up_bias, np.float32), opt(gate_bias, np.float32)
        p = lambda a: None if a is None else _ptr(a)
        out = np.zeros((N, M), dtype=np.float32)
        rc = self.lib.refh_mul_mat_glu(reg, dev, n_threads, t, M, K, N, _ptr(W_up), p(W_gate), _ptr(X), p(up_bias), p(gate_bias), glu_op,
                                       alpha, limit, _ptr(out))
        if rc:
            raise RuntimeError(f"refh_mul_mat_glu rc={rc} (2 = backend does not support a node)")
        return out

    def mul_mat_id_glu(self, reg: int, t: int, W_up, W_gate, B, ids, up_bias=None, gate_bias=None, glu_op: int = -1, alpha=1.702, limit=7.0,
                       n_threads=0, dev=0) -> np.ndarray:
        """the MUL_MAT_ID / ADD_ID forms: W uint8 [n_as, M, row_bytes]; B [n_tok, ne11, K]; ids [n_tok, n_used]; biases [n_as, M]"""
        B = np.ascontiguousarray(B, dtype=np.float32)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        W_up = np.ascontiguousarray(W_up, dtype=np.uint8)
        n_tok, ne11, K = B.shape
        n_as, M, _ = W_up.shape
        n_used = ids.shape[1]
        opt = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)
        W_gate, up_bias, gate_bias = opt(W_gate, np.uint8), opt(up_bias, np.float32), opt(gate_bias, np.float32)
        p = lambda a: None if a is None else _ptr(a)
        out = np.zeros((n_tok, n_used, M), dtype=np.float32)
        rc = self.lib.refh_mul_mat_id_glu(reg, dev, n_threads, t, M, K, n_as, n_used, ne11, n_tok, _ptr(W_up), p(W_gate), _ptr(B), _ptr(ids),
                                          p(up_bias), p(gate_bias), glu_op, alpha, limit, _ptr(out))
        if rc:
            raise RuntimeError(f"refh_mul_mat_id_glu rc={rc} (2 = backend does not support a node)")
        return out

    def mul_mat_stack(self, reg: int, t: int, shapes, N: int, W_all: np.ndarray, X_all: np.ndarray, n_threads=0,
                      dev=0, seconds=1.0, reps=1):
        """times the stack of MUL_MATs as ONE graph holding `reps` repetitions of it; returns (us per repetition, repetitions run)"""
        Ms = np.array([s[0] for s in shapes], dtype=np.int64)
        Ks = np.array([s[1] for s in shapes], dtype=np.int64)
        us = C.c_double(0)
        runs = C.c_int(0)
        rc = self.lib.refh_mul_mat_stack(reg, dev, n_threads, t, len(shapes), _ptr(Ms), _ptr(Ks), N,
                                         _ptr(W_all), _ptr(X_all), seconds, int(reps), C.byref(us), C.byref(runs))
        if rc:
            raise RuntimeError(f"refh_mul_mat_stack rc={rc}")
        return us.value, runs.value


def have_ref() -> bool:
    return all((REF_DIR / n).exists() for n in ("libggml-base.so", "libggml.so", "libref_harness.so", "libggml-cpu-x64.so"))


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1
