"""K7: GLU / bias epilogues fused into the one-column mat-vec, and the stand-alone GLU / ADD / ADD_ID nodes.
Reference behaviour: ggml_cuda_should_fuse_mul_mat + the graph loop (ggml-cuda.cu:2020-2176, 3404-3600), mul_mat_vec_q's epilogue
(mmvq.cu:181-328), test_mul_mat_vec_fusion (tests/test-backend-ops.cpp:5120-5249, NMSE bar 5e-3 there).

CPU part (-m "not gpu"): the oracle's node functions against the committed fixture tests/golden/fusion_k7.npz (whole graphs run on
the reference's ggml-cpu) and, when oracle/_ref is present, against the compiled reference on fresh inputs.
GPU part: the C-ABI against the oracle and against the fixture."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from oracle.pyoracle import Q4_0, Q4_K, Q6_K, Q8_0, TYPE_NAMES, have_ref

GOLD = Path(__file__).resolve().parent / "golden"
# ggml-cpu evaluates GEGLU / GEGLU_QUICK through fp16 tables and SWIGLU through a polynomial expf (vec.h:1326-1340, vec.cpp:376-410);
# the float formulas agree with them to ~7e-8 NMSE. REGLU, SWIGLU_OAI, GEGLU_ERF are the same scalar code: bit-identical.
NMSE_REF = 1e-6
NMSE_TIGHT = 1e-10   # GPU vs oracle: identical integer block sums, libm-accurate device expf / tanhf / erff
EXACT_OPS = (0, 3, 4)
GLU_NAMES = {0: "reglu", 1: "geglu", 2: "swiglu", 3: "swiglu_oai", 4: "geglu_erf", 5: "geglu_quick"}


# ---------------------------------------------------------------- oracle vs the reference (CPU)

@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q6_K])
def test_oracle_fused_patterns_vs_golden(oracle, t):
    g = np.load(GOLD / "fusion_k7.npz")
    tn = TYPE_NAMES[t]
    Wu, Wg, X, bu, bg = (g[f"{tn}_{k}"] for k in ("Wu", "Wg", "X", "bu", "bg"))
    for op in range(6):
        for bias in (False, True):
            ref = g[f"{tn}_glu{op}" + ("_bias" if bias else "")]
            got = oracle.mul_mat_glu(t, Wu, Wg, X, bu if bias else None, bg if bias else None, glu_op=op)
            if op in EXACT_OPS:
                assert np.array_equal(ref, got), (tn, GLU_NAMES[op], bias)
            else:
                assert oracle.nmse(ref, got) <= NMSE_REF, (tn, GLU_NAMES[op], bias)
    assert np.array_equal(g[f"{tn}_bias_only"], oracle.mul_mat_glu(t, Wu, None, X, bu, None, glu_op=-1))


def test_oracle_fused_mul_mat_id_vs_golden(oracle):
    g = np.load(GOLD / "fusion_k7.npz")
    Wu, Wg, B, ids, bu, bg = (g[f"id_{k}"] for k in ("Wu", "Wg", "B", "ids", "bu", "bg"))
    for op in (1, 2, 3):
        for bias in (False, True):
            ref = g[f"id_glu{op}" + ("_bias" if bias else "")]
            got = oracle.mul_mat_id_glu(Q4_K, Wu, Wg, B, ids, bu if bias else None, bg if bias else None, glu_op=op)
            assert oracle.nmse(ref, got) <= (0.0 if op in EXACT_OPS else NMSE_REF), (GLU_NAMES[op], bias)
    assert np.array_equal(g["id_bias_only"], oracle.mul_mat_id_glu(Q4_K, Wu, None, B, ids, bu, None, glu_op=-1))


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
def test_oracle_glu_vs_compiled_reference(oracle, harness):
    """fresh inputs incl. several columns (broadcast ADD of the bias row) and the extremes of the activations"""
    from oracle.pyoracle import RefQuant
    rng = np.random.default_rng(77)
    cpu = harness.cpu("x64")
    M, K, N = 40, 256, 3
    Wu = RefQuant().quantize_weights(Q8_0, rng.uniform(-4, 4, (M, K)).astype(np.float32))
    Wg = RefQuant().quantize_weights(Q8_0, rng.uniform(-4, 4, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)   # dot products reach +-40: beyond the +-10 table range and SWIGLU_OAI's limit
    bu, bg = rng.uniform(-1, 1, M).astype(np.float32), rng.uniform(-1, 1, M).astype(np.float32)
    for op in range(6):
        ref = harness.mul_mat_glu(cpu, Q8_0, Wu, Wg, X, bu, bg, glu_op=op, alpha=1.3, limit=5.0, n_threads=1)
        got = oracle.mul_mat_glu(Q8_0, Wu, Wg, X, bu, bg, glu_op=op, alpha=1.3, limit=5.0)
        assert oracle.nmse(ref, got) <= (0.0 if op in EXACT_OPS else NMSE_REF), GLU_NAMES[op]


# ---------------------------------------------------------------- the C-ABI on the GPU

def cu(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q6_K])
def test_fused_mul_mat_golden(ctx, oracle, pkg, t):
    g = np.load(GOLD / "fusion_k7.npz")
    tn = TYPE_NAMES[t]
    Wu, Wg, X, bu, bg = (g[f"{tn}_{k}"] for k in ("Wu", "Wg", "X", "bu", "bg"))
    if t == Q6_K:  # the fixture's K = 512 is not a TMA-staged q6_K shape (K % 2048): the fused call must decline, not guess
        with pytest.raises(pkg.B200Error):
            ctx.mul_mat_fused(t, cu(Wu), cu(Wg), cu(X), glu_op=2)
        return
    for op in range(6):
        for bias in (False, True):
            ref = g[f"{tn}_glu{op}" + ("_bias" if bias else "")]
            l0 = ctx.launches
            got = ctx.mul_mat_fused(t, cu(Wu), cu(Wg), cu(X), cu(bu) if bias else None, cu(bg) if bias else None, glu_op=op).cpu().numpy()
            assert ctx.launches - l0 == 1, "the fused pattern is ONE launch"
            assert oracle.nmse(ref, got) <= NMSE_REF, (tn, GLU_NAMES[op], bias)
            orc = oracle.mul_mat_glu(t, Wu, Wg, X, bu if bias else None, bg if bias else None, glu_op=op)
            assert oracle.nmse(orc, got) <= NMSE_TIGHT, (tn, GLU_NAMES[op], bias)
    got = ctx.mul_mat_fused(t, cu(Wu), None, cu(X), cu(bu), None, glu_op=-1).cpu().numpy()
    assert oracle.nmse(g[f"{tn}_bias_only"], got) <= NMSE_TIGHT


@pytest.mark.gpu
@pytest.mark.parametrize("t,M,K", [(Q4_K, 1000, 4096), (Q4_0, 333, 2048), (Q8_0, 64, 8192 + 256), (Q6_K, 130, 2048), (Q4_K, 77, 14336),
                                   (Q4_K, 5, 256), (Q6_K, 48, 6144), (Q4_0, 20, 28672)])
def test_fused_mul_mat_shapes(ctx, oracle, t, M, K):
    """row groups of 1..32 rows (U <= 32), rows longer than one ring stage (U > 32), ragged M, the pre-quantized activation path
    (K = 28672: 112 units > the fused quantizer's bound)"""
    from test_gpu_parity import quantize_weights
    rng = np.random.default_rng(M * 7 + K)
    Wu = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    Wg = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.uniform(-1, 1, (1, K)) * 0.2).astype(np.float32)
    bu, bg = rng.uniform(-1, 1, M).astype(np.float32), rng.uniform(-1, 1, M).astype(np.float32)
    for op, bias in ((2, False), (2, True), (1, True), (3, False), (0, True), (4, False), (5, True)):
        got = ctx.mul_mat_fused(t, cu(Wu), cu(Wg), cu(X), cu(bu) if bias else None, cu(bg) if bias else None, glu_op=op, alpha=1.5, limit=3.0)
        orc = oracle.mul_mat_glu(t, Wu, Wg, X, bu if bias else None, bg if bias else None, glu_op=op, alpha=1.5, limit=3.0)
        assert oracle.nmse(orc, got.cpu().numpy()) <= NMSE_TIGHT, (GLU_NAMES[op], bias)
    got = ctx.mul_mat_fused(t, cu(Wu), None, cu(X), cu(bu), None, glu_op=-1)
    assert oracle.nmse(oracle.mul_mat_glu(t, Wu, None, X, bu, None, -1), got.cpu().numpy()) <= NMSE_TIGHT
    # the unfused composition through the same C-ABI gives the same numbers (same integer sums, same epilogue arithmetic)
    up = ctx.add(ctx.mul_mat(t, cu(Wu), cu(X)), cu(bu))
    gate = ctx.add(ctx.mul_mat(t, cu(Wg), cu(X)), cu(bg))
    unfused = ctx.glu(2, gate, up).cpu().numpy()
    fused = ctx.mul_mat_fused(t, cu(Wu), cu(Wg), cu(X), cu(bu), cu(bg), glu_op=2).cpu().numpy()
    assert oracle.nmse(unfused, fused) <= NMSE_TIGHT


@pytest.mark.gpu
def test_fused_mul_mat_batched(ctx, oracle, pkg):
    """test_mul_mat_vec_fusion's layout: 4-D weights [K, M, 4, 2], one column per batch, bias of the output's shape"""
    import torch
    from test_gpu_parity import quantize_weights
    rng = np.random.default_rng(5)
    t, M, K, ne2, ne3 = Q4_K, 40, 512, 4, 2
    Wu = quantize_weights(oracle, t, rng.uniform(-1, 1, (ne3 * ne2 * M, K)).astype(np.float32)).reshape(ne3, ne2, M, -1)
    Wg = quantize_weights(oracle, t, rng.uniform(-1, 1, (ne3 * ne2 * M, K)).astype(np.float32)).reshape(ne3, ne2, M, -1)
    X = rng.uniform(-1, 1, (ne3, ne2, 1, K)).astype(np.float32)
    bu = rng.uniform(-1, 1, (ne3, ne2, 1, M)).astype(np.float32)
    bg = rng.uniform(-1, 1, (ne3, ne2, 1, M)).astype(np.float32)
    desc = pkg.MulMatDesc.contiguous(t, M, K, 1, ne2, ne3)
    dbu, dbg = cu(bu), cu(bg)
    f = pkg.Fusion(2, 0.0, 0.0, dbu.data_ptr(), dbg.data_ptr(), 0, 4 * M, 4 * M * ne2)
    out = torch.empty((ne3, ne2, 1, M), dtype=torch.float32, device="cuda")
    dWu, dWg, dX = cu(Wu), cu(Wg), cu(X)
    pkg.check(pkg.lib().b200_mul_mat_fused(ctx.h, C.byref(desc), dWu.data_ptr(), dWg.data_ptr(), dX.data_ptr(), C.byref(f), out.data_ptr()))
    up = oracle.mul_mat(t, Wu, X, ne2, ne3) + bu
    gate = oracle.mul_mat(t, Wg, X, ne2, ne3) + bg
    assert oracle.nmse(oracle.glu(2, gate, up), out.cpu().numpy()) <= NMSE_TIGHT


@pytest.mark.gpu
def test_fused_mul_mat_id_golden_and_shapes(ctx, oracle):
    from test_gpu_parity import quantize_weights
    g = np.load(GOLD / "fusion_k7.npz")
    Wu, Wg, B, ids, bu, bg = (g[f"id_{k}"] for k in ("Wu", "Wg", "B", "ids", "bu", "bg"))
    for op in (1, 2, 3):
        for bias in (False, True):
            ref = g[f"id_glu{op}" + ("_bias" if bias else "")]
            l0 = ctx.launches
            got = ctx.mul_mat_id_fused(Q4_K, cu(Wu), cu(Wg), cu(B), cu(ids), cu(bu) if bias else None, cu(bg) if bias else None, glu_op=op)
            assert ctx.launches - l0 == 1
            assert oracle.nmse(ref, got.cpu().numpy()) <= NMSE_REF, (GLU_NAMES[op], bias)
    got = ctx.mul_mat_id_fused(Q4_K, cu(Wu), None, cu(B), cu(ids), cu(bu), None, glu_op=-1)
    assert oracle.nmse(g["id_bias_only"], got.cpu().numpy()) <= NMSE_TIGHT
    rng = np.random.default_rng(9)
    for t, n_as, M, K, n_used, ne11 in ((Q4_0, 8, 200, 4096, 4, 1), (Q8_0, 3, 33, 512, 2, 2), (Q6_K, 4, 64, 2048, 3, 1), (Q4_K, 16, 96, 12288, 8, 1)):
        Wu = quantize_weights(oracle, t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
        Wg = quantize_weights(oracle, t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
        B = (rng.uniform(-1, 1, (1, ne11, K)) * 0.2).astype(np.float32)
        ids = rng.permutation(n_as)[:n_used].astype(np.int32).reshape(1, n_used)
        bu, bg = rng.uniform(-1, 1, (n_as, M)).astype(np.float32), rng.uniform(-1, 1, (n_as, M)).astype(np.float32)
        for op, bias in ((2, True), (3, False), (1, True)):
            got = ctx.mul_mat_id_fused(t, cu(Wu), cu(Wg), cu(B), cu(ids), cu(bu) if bias else None, cu(bg) if bias else None, glu_op=op)
            orc = oracle.mul_mat_id_glu(t, Wu, Wg, B, ids, bu if bias else None, bg if bias else None, glu_op=op)
            assert oracle.nmse(orc, got.cpu().numpy()) <= NMSE_TIGHT, (TYPE_NAMES[t], GLU_NAMES[op], bias)


@pytest.mark.gpu
def test_standalone_glu_add_add_id(ctx, oracle):
    rng = np.random.default_rng(11)
    gate = (rng.standard_normal((7, 333)) * 6).astype(np.float32)
    up = rng.standard_normal((7, 333)).astype(np.float32)
    for op in range(6):
        got = ctx.glu(op, cu(gate), cu(up), alpha=1.702, limit=7.0).cpu().numpy()
        ref = oracle.glu(op, gate, up)
        # device tanhf / erff / expf are within a few ulp of glibc's; 1 + tanh(..) cancels for large negative gates
        assert oracle.nmse(ref, got) <= 1e-12 and np.allclose(got, ref, rtol=1e-4, atol=1e-6), GLU_NAMES[op]
    # the single-tensor form (halves of one row) is the caller's pointer arithmetic: row stride 2 * nc
    both = cu(np.concatenate([gate, up], axis=1))
    got = ctx.glu(2, both[:, :333], both[:, 333:]).cpu().numpy()
    assert np.allclose(got, oracle.glu(2, gate, up), rtol=2e-6, atol=1e-7)
    # ADD with ggml broadcasting: same shape, a bias row, a repeated column block, 4-D
    a = rng.standard_normal((2, 3, 5, 64)).astype(np.float32)
    for bshape in ((2, 3, 5, 64), (1, 1, 1, 64), (1, 3, 1, 64), (2, 1, 5, 32), (1, 1, 5, 1)):
        b = rng.standard_normal(bshape).astype(np.float32)
        ref = a + np.tile(b, [a.shape[i] // b.shape[i] for i in range(4)])
        assert np.array_equal(ctx.add(cu(a), cu(b)).cpu().numpy(), ref), bshape
        assert np.array_equal(ctx.mul(cu(a), cu(b)).cpu().numpy(), a * np.tile(b, [a.shape[i] // b.shape[i] for i in range(4)])), bshape
    # ADD_ID
    n_tok, n_used, M, n_as = 6, 3, 50, 5
    x = rng.standard_normal((n_tok, n_used, M)).astype(np.float32)
    bias = rng.standard_normal((n_as, M)).astype(np.float32)
    ids = rng.integers(0, n_as, (n_tok, n_used)).astype(np.int32)
    assert np.array_equal(ctx.add_id(cu(x), cu(bias), cu(ids)).cpu().numpy(), oracle.add_id(x, bias, ids))
