"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/ggml_b200.h declares,
answers geometry queries without a GPU, and FAILS LOUDLY (no CPU fallback) when asked to compute without one."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    hdr = (ROOT / "include" / "ggml_b200.h").read_text()
    return sorted(set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", hdr)))


def test_header_symbols_exported(pkg):
    syms = declared_symbols()
    assert len(syms) >= 40
    L = pkg.lib()
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing
    assert sorted(pkg.EXPORTS) == syms


def test_backend_library_exports_ggml_entry_points(pkg):
    if not pkg.BACKEND_PATH.exists():
        pytest.skip("libggml-b200.so not built (needs the reference headers)")
    out = subprocess.run(["nm", "-D", "--defined-only", str(pkg.BACKEND_PATH)], capture_output=True, text=True, check=True).stdout
    for s in ("ggml_backend_init", "ggml_backend_score", "ggml_backend_b200_reg"):
        assert re.search(rf"\bT {s}\b", out), s


def test_supported_geometry_queries(pkg):
    L = pkg.lib()
    d = pkg.MulMatDesc.contiguous(pkg.Q4_0, 4096, 4096, 1)
    assert L.b200_mul_mat_supported(C.byref(d)) == 1
    d = pkg.MulMatDesc.contiguous(pkg.Q4_K, 128, 512, 3, 2, 1, 4, 2)
    assert L.b200_mul_mat_supported(C.byref(d)) == 1
    d = pkg.MulMatDesc.contiguous(pkg.Q4_0, 64, 64, 1)
    d.type = 10  # q2_K is outside the hot path
    assert L.b200_mul_mat_supported(C.byref(d)) == 0
    d = pkg.MulMatDesc.contiguous(pkg.Q4_K, 64, 256, 1)
    d.ne00 = 128  # not a multiple of the 256 super-block
    assert L.b200_mul_mat_supported(C.byref(d)) == 0
    d = pkg.MulMatDesc.contiguous(pkg.Q8_0, 64, 256, 1, 3, 1, 4, 1)  # 4 % 3 != 0: not broadcastable
    assert L.b200_mul_mat_supported(C.byref(d)) == 0
    e = pkg.MulMatIdDesc.contiguous(pkg.Q4_0, 512, 256, 8, 2, 1, 17)
    assert L.b200_mul_mat_id_supported(C.byref(e)) == 1


def _t(pkg, gtype, ne, nb=None, esize=None):
    """a b200_tensor descriptor without memory behind it (the *_supported predicates only look at geometry): contiguous unless nb is given"""
    esize = esize or {0: 4, 1: 2, 26: 4, 27: 8}[gtype]
    ne = list(ne) + [1] * (4 - len(ne))
    if nb is None:
        nb = [esize]
        for i in range(3):
            nb.append(nb[-1] * ne[i])
    return pkg.Tensor(0x1000, gtype, (C.c_int64 * 4)(*ne), (C.c_size_t * 4)(*nb))


def test_f3_supported_predicates(pkg):
    """what dev_supports_op answers for the f3 ops is decided by these host-side predicates: the Llama decode shapes
    are accepted, what the kernels do not implement is declined (and then stays on ggml-cpu)"""
    L = pkg.lib()
    F32, F16, I32, I64 = 0, 1, 26, 27
    x = _t(pkg, F32, [4096, 7])
    assert L.b200_rms_norm_supported(x, x) == 1
    assert L.b200_rms_norm_supported(_t(pkg, F16, [4096, 7]), x) == 0
    assert L.b200_rms_norm_mul_supported(x, _t(pkg, F32, [4096]), x) == 1
    assert L.b200_rms_norm_mul_supported(x, _t(pkg, F32, [1]), x) == 0              # a scalar weight is not the per-channel MUL
    # GET_ROWS: quantized token embedding by I32 ids; I64 ids and F16 destinations are not ggml's GET_ROWS
    emb = pkg.Tensor(0x1000, pkg.Q4_K, (C.c_int64 * 4)(4096, 32000, 1, 1), (C.c_size_t * 4)(144, 2304, 2304 * 32000, 2304 * 32000))
    assert L.b200_get_rows_supported(emb, _t(pkg, I32, [5]), _t(pkg, F32, [4096, 5])) == 1
    assert L.b200_get_rows_supported(emb, _t(pkg, I64, [5]), _t(pkg, F32, [4096, 5])) == 0
    assert L.b200_get_rows_supported(emb, _t(pkg, I32, [5]), _t(pkg, F32, [4096, 6])) == 0
    # ROPE: normal and neox over part or all of the head; mrope / vision modes stay on ggml-cpu
    q = _t(pkg, F32, [128, 32, 3])
    pos = _t(pkg, I32, [3])
    rp = pkg.RopeParams(128, 0, 8192, 500000.0, 1.0, 0.0, 1.0, 32.0, 1.0)
    assert L.b200_rope_supported(q, pos, None, q, C.byref(rp)) == 1
    rp.mode = 2
    assert L.b200_rope_supported(q, pos, None, q, C.byref(rp)) == 1
    rp.mode = 8
    assert L.b200_rope_supported(q, pos, None, q, C.byref(rp)) == 0
    rp.mode, rp.n_dims = 0, 130
    assert L.b200_rope_supported(q, pos, None, q, C.byref(rp)) == 0
    # SET_ROWS: F32 rows into the F16 KV cache by I64 slot index
    kc = _t(pkg, F16, [1024, 4096])
    assert L.b200_set_rows_supported(_t(pkg, F32, [1024, 3]), _t(pkg, I64, [3]), kc) == 1
    assert L.b200_set_rows_supported(_t(pkg, F32, [1024, 3]), _t(pkg, I64, [4]), kc) == 0
    assert L.b200_set_rows_supported(_t(pkg, F16, [1024, 3]), _t(pkg, I64, [3]), kc) == 0
    # CPY: F32 / F16 in any combination, equal element counts
    assert L.b200_cpy_supported(_t(pkg, F32, [64, 8]), _t(pkg, F16, [512])) == 1
    assert L.b200_cpy_supported(_t(pkg, F32, [64, 8]), _t(pkg, F16, [511])) == 0
    assert L.b200_cpy_supported(_t(pkg, I32, [64]), _t(pkg, F32, [64])) == 0
    # FLASH_ATTN_EXT: F32 q, F16 K/V views of the cache (heads interleaved per position), F16 mask; head sizes 64 / 128 / 256
    def fa(D, nh, nkv, nq, n_kv, ktype=F16):
        qt = _t(pkg, F32, [D, nq, nh])
        kt = _t(pkg, ktype, [D, n_kv, nkv], nb=[2, 2 * D * nkv, 2 * D, 2 * D * nkv * n_kv], esize=2)
        m = _t(pkg, F16, [n_kv, (nq + 63) // 64 * 64])
        return L.b200_flash_attn_ext_supported(qt, kt, kt, m, None, _t(pkg, F32, [D, nh, nq]))
    assert fa(128, 32, 8, 1, 512) == 1 and fa(64, 8, 8, 300, 1000) == 1 and fa(256, 8, 2, 1, 77) == 1
    assert fa(80, 8, 8, 1, 64) == 0 and fa(128, 32, 5, 1, 64) == 0     # head size 80; 32 heads over 5 KV heads
    assert fa(128, 8, 8, 1, 64, ktype=F32) == 0                          # F32 (or quantized) K/V caches stay on ggml-cpu


def test_python_structs_match_the_header(pkg, tmp_path):
    """the ctypes mirrors of the C-ABI structs have the size and field offsets the C compiler gives include/ggml_b200.h"""
    structs = {"b200_mul_mat_desc": pkg.MulMatDesc, "b200_mul_mat_id_desc": pkg.MulMatIdDesc, "b200_row_split": pkg.RowSplit,
               "b200_row_split_ll": pkg.RowSplitLL, "b200_fusion": pkg.Fusion, "b200_tensor": pkg.Tensor, "b200_rope_params": pkg.RopeParams,
               "b200_norm": pkg.Norm}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "ggml_b200.h"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append(f'  printf("{cname} %zu", sizeof({cname}));')
        for fname, *_ in cls._fields_:
            lines.append(f'  printf(" %zu", offsetof({cname}, {fname}));')
        lines.append('  printf("\\n");')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-I", str(ROOT / "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines()
    for line, (cname, cls) in zip(out, structs.items()):
        parts = line.split()
        assert parts[0] == cname
        assert int(parts[1]) == C.sizeof(cls), cname
        assert [int(x) for x in parts[2:]] == [getattr(cls, f[0]).offset for f in cls._fields_], cname


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = pkg.lib()
    assert L.b200_device_count() == 0
    h = C.c_void_p()
    rc = L.b200_ctx_create(0, None, C.byref(h))
    assert rc != 0 and not h.value
    assert b"no CPU fallback" in L.b200_last_error()
    with pytest.raises(pkg.B200Error):
        pkg.Context(0)


def test_backend_registers_no_device_without_gpu(pkg, harness):
    import torch
    if torch.cuda.is_available() or not pkg.BACKEND_PATH.exists():
        pytest.skip("GPU present or backend not built")
    # ggml_backend_score() == 0 -> the registry refuses to load it (ggml-backend-reg.cpp:264-270)
    with pytest.raises(RuntimeError):
        harness.load(pkg.BACKEND_PATH)
