"""CPU tests: pin oracle/oracle.c against (a) the committed golden fixtures generated from the reference's own compiled
code and (b) that compiled code itself when oracle/_ref is present (differential, seeded)."""
import glob
from pathlib import Path

import numpy as np
import pytest

from oracle.pyoracle import Q4_0, Q4_K, Q5_K, Q6_K, Q8_0, TYPE_NAMES, row_size

GOLD = Path(__file__).resolve().parent / "golden"


def mask_q8k_zero_blocks(a: np.ndarray) -> np.ndarray:
    """bsums of an all-zero q8_K block are not written by quantize_row_q8_K_ref (ggml-quants.c:2570-2575): ignore them"""
    a = a.copy().reshape(a.shape[0], -1, 292)
    zero = (a[:, :, :4].view(np.float32)[:, :, 0] == 0)
    a[zero, 260:] = 0
    return a


def test_act_quant_golden(oracle):
    g = np.load(GOLD / "act_quant.npz")
    for name in ("q8_0", "q8_1", "q8_K"):
        mine, ref = oracle.quantize(name, g["x"]), g[name]
        if name == "q8_K":
            mine, ref = mask_q8k_zero_blocks(mine), mask_q8k_zero_blocks(ref)
        assert np.array_equal(mine, ref), name


@pytest.mark.parametrize("f", sorted(glob.glob(str(GOLD / "mul_mat_q*_*x*.npz"))))
def test_mul_mat_golden(oracle, f):
    g = np.load(f)
    t = int(g["type"])
    out = oracle.mul_mat(t, g["W"], g["X"])
    # the oracle restates the generic scalar code: bit-for-bit equal to ggml-cpu built without SIMD
    assert np.array_equal(out.view(np.uint32), g["out"].view(np.uint32))
    if t == Q4_0 and g["w_f32"].size:
        assert np.array_equal(oracle.quantize("q4_0", g["w_f32"]), g["W"])


def test_mul_mat_bcast_golden(oracle):
    g = np.load(GOLD / "mul_mat_q4_K_bcast.npz")
    out = oracle.mul_mat(int(g["type"]), g["W"], g["X"], ne02=int(g["ne02"]), ne03=int(g["ne03"]))
    assert np.array_equal(out.view(np.uint32), g["out"].view(np.uint32))


@pytest.mark.parametrize("f", sorted(glob.glob(str(GOLD / "mul_mat_id_*.npz"))))
def test_mul_mat_id_golden(oracle, f):
    g = np.load(f)
    out = oracle.mul_mat_id(int(g["type"]), g["W"], g["B"], g["ids"])
    assert np.array_equal(out.view(np.uint32), g["out"].view(np.uint32))


def test_fp16_conversions(oracle):
    hs = np.arange(65536, dtype=np.uint16)
    f = hs.view(np.float16).astype(np.float32)
    ok = ~np.isnan(f)
    mine = np.array([oracle.lib.orc_fp16_to_fp32(int(v)) for v in hs], dtype=np.float32)
    assert np.array_equal(mine.view(np.uint32)[ok], f.view(np.uint32)[ok])
    rng = np.random.default_rng(1)
    x = (rng.standard_normal(50000) * 10.0 ** rng.integers(-9, 6, 50000)).astype(np.float32)
    x = np.concatenate([x, f[ok], np.array([65519.9, 65520.0, 5.96e-8, 2.98e-8, 2.9802322e-8, 0.0, -0.0], dtype=np.float32)])
    with np.errstate(over="ignore"):
        want = x.astype(np.float16).view(np.uint16)
    got = np.array([oracle.lib.orc_fp32_to_fp16(float(v)) for v in x], dtype=np.uint16)
    assert np.array_equal(want, got)


# ---- differential against the reference's compiled code (skipped where oracle/_ref is absent)

@pytest.mark.parametrize("seed", [0, 1, 2])
def test_quantizers_vs_ref(oracle, refq, seed):
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal((32, 2048)) * rng.choice([1e-3, 1.0, 50.0], (32, 1))).astype(np.float32)
    x[rng.integers(0, 32), :512] = 0
    for name in ("q8_0", "q8_1", "q8_K", "q4_0"):
        a, b = oracle.quantize(name, x), refq.quantize_ref(name, x)
        if name == "q8_K":
            a, b = mask_q8k_zero_blocks(a), mask_q8k_zero_blocks(b)
        assert np.array_equal(a, b), name


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
def test_dequant_and_mul_mat_vs_ref(oracle, refq, harness, t):
    rng = np.random.default_rng(100 + t)
    M, K, N = 40, 768, 7
    W = refq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    assert np.array_equal(oracle.dequantize(t, W, K), refq.dequantize(t, W, K))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    mine = oracle.mul_mat(t, W, X)
    strict, _ = harness.mul_mat(harness.cpu("x64"), t, W, X, n_threads=2)
    assert np.array_equal(mine.view(np.uint32), strict.view(np.uint32))
    # the SIMD builds (AVX2 / AVX512-VNNI + tinyBLAS) differ only by fp32 summation order and the a8 rounding ties
    fast, _ = harness.mul_mat(harness.cpu(harness.best_cpu_variant()), t, W, X, n_threads=4)
    assert oracle.nmse(fast, mine) < 1e-9


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
def test_mul_mat_id_vs_ref(oracle, refq, harness, t):
    rng = np.random.default_rng(200 + t)
    n_as, M, K, n_used, n_tok = 8, 24, 512, 4, 9
    W = refq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    mine = oracle.mul_mat_id(t, W, B, ids)
    ref, _ = harness.mul_mat_id(harness.cpu("x64"), t, W, B, ids, n_threads=2)
    assert np.array_equal(mine.view(np.uint32), ref.view(np.uint32))


def test_nmse_matches_definition(oracle):
    a = np.array([1.0, 2.0, 3.0], dtype=np.float32)
    b = np.array([1.0, 2.5, 3.0], dtype=np.float32)
    assert abs(oracle.nmse(a, b) - 0.25 / 14.0) < 1e-12
