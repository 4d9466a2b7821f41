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


def test_python_structs_match_the_header(pkg, tmp_path):
    """the ctypes mirrors of the C-ABI structs have the size and field offsets the C compiler gives include/ggml_b200.h"""
    structs = {"b200_mul_mat_desc": pkg.MulMatDesc, "b200_mul_mat_id_desc": pkg.MulMatIdDesc, "b200_row_split": pkg.RowSplit,
               "b200_row_split_ll": pkg.RowSplitLL, "b200_fusion": pkg.Fusion}
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
