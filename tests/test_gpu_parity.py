"""GPU parity tests (pytest -m gpu): every call goes through the C-ABI (libb200mm.so via ctypes); the oracle is only the checker.
Bar: activation quantizer bit-exact; MUL_MAT / MUL_MAT_ID NMSE <= 5e-4 against ggml-cpu semantics
(tests/test-backend-ops.cpp:3569-3571, 3697-3699 in the reference) — in practice ~1e-13 because the integer block sums are identical."""
import glob
from pathlib import Path

import numpy as np
import pytest

from oracle.pyoracle import Q4_0, Q4_K, Q5_K, Q6_K, Q8_0, TYPE_NAMES, have_ref

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"
NMSE_MAX = 5e-4          # the reference's own threshold
NMSE_TIGHT = 1e-10       # what identical integer sums + fp32 reordering actually give
NMSE_F16 = 1e-6          # prefill on kind::f16 (default): operands rounded to 11 significant bits, fp32 accumulation; measured 1.2e-7


def cu(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def quantize_weights(oracle, t, w):
    if have_ref():
        from oracle.pyoracle import RefQuant
        return RefQuant().quantize_weights(t, w)
    if t in (Q4_K, Q5_K, Q6_K):
        pytest.skip("K-quant weights need oracle/_ref")
    return oracle.quantize("q4_0" if t == Q4_0 else "q8_0", w)


def mask_q8k(a):
    a = a.copy().reshape(a.shape[0], -1, 292)
    zero = (a[:, :, :4].view(np.float32)[:, :, 0] == 0)
    a[zero, 260:] = 0
    return a


# ---------------------------------------------------------------- K1

def test_act_quant_golden_bit_exact(ctx):
    g = np.load(GOLD / "act_quant.npz")
    x = cu(g["x"])
    for name in ("q8_0", "q8_1", "q8_K"):
        got = ctx.quantize(name, x).cpu().numpy()
        ref = g[name]
        if name == "q8_K":
            got, ref = mask_q8k(got), mask_q8k(ref)
        assert np.array_equal(got, ref), name


@pytest.mark.parametrize("k", [32, 96, 256, 4096, 14336])
def test_act_quant_vs_oracle(ctx, oracle, k):
    rng = np.random.default_rng(k)
    x = (rng.standard_normal((9, k)) * rng.choice([1e-4, 1.0, 300.0], (9, 1))).astype(np.float32)
    x[0, :32] = 0
    for name in ("q8_0", "q8_1") + (("q8_K",) if k % 256 == 0 else ()):
        got = ctx.quantize(name, cu(x)).cpu().numpy()
        ref = oracle.quantize(name, x)
        if name == "q8_K":
            got, ref = mask_q8k(got), mask_q8k(ref)
        assert np.array_equal(got, ref), (name, k)


def test_act_quant_strided_rows(ctx, oracle):
    import torch
    rng = np.random.default_rng(5)
    big = rng.uniform(-1, 1, (6, 1024 + 36)).astype(np.float32)
    xb = cu(big)
    view = xb[:, 4:4 + 1024]                       # row stride 1060 floats, base offset 16 B... and
    got = ctx.quantize("q8_0", view).cpu().numpy()
    assert np.array_equal(got, oracle.quantize("q8_0", big[:, 4:4 + 1024]))
    view = xb[:, 1:1 + 1024]                       # only 4-byte aligned: scalar-load path
    got = ctx.quantize("q8_K", view).cpu().numpy()
    assert np.array_equal(mask_q8k(got), mask_q8k(oracle.quantize("q8_K", big[:, 1:1 + 1024])))


def test_unit_records_hold_the_same_values(ctx, oracle):
    rng = np.random.default_rng(11)
    x = rng.uniform(-2, 2, (5, 1024)).astype(np.float32)
    # kind 0: q8_0 values
    main, aux = ctx.quantize_units(0, cu(x))
    main, aux = main.cpu().numpy(), aux.cpu().numpy()
    ref = oracle.quantize("q8_0", x).reshape(5, -1, 34)
    q = main[:, :, :256].reshape(5, -1, 32)
    assert np.array_equal(q, ref[:, :, 2:])
    assert np.array_equal(main[:, :, 256:].reshape(5, -1, 2), ref[:, :, :2])
    sums = aux.view(np.int16).reshape(5, -1)
    assert np.array_equal(sums, ref[:, :, 2:].view(np.int8).astype(np.int32).sum(-1).astype(np.int16))
    # kind 1: q8_K values
    main, aux = ctx.quantize_units(1, cu(x))
    main, aux = main.cpu().numpy(), aux.cpu().numpy()
    ref = oracle.quantize("q8_K", x).reshape(5, -1, 292)
    assert np.array_equal(main[:, :, :256], ref[:, :, 4:260])
    assert np.array_equal(aux[:, :, :4], ref[:, :, :4])
    bs16 = ref[:, :, 260:].view(np.int16).astype(np.int32).reshape(5, -1, 8, 2).sum(-1).astype(np.int16)
    assert np.array_equal(main[:, :, 256:].view(np.int16), bs16)


# ---------------------------------------------------------------- K3 on golden fixtures

@pytest.mark.parametrize("f", sorted(glob.glob(str(GOLD / "mul_mat_q*_*x*.npz"))))
def test_mul_mat_golden(ctx, oracle, f):
    g = np.load(f)
    t = int(g["type"])
    got = ctx.mul_mat(t, cu(g["W"]), cu(g["X"])).cpu().numpy()
    assert oracle.nmse(g["out"], got) <= NMSE_TIGHT


def test_mul_mat_bcast_golden(ctx, oracle):
    g = np.load(GOLD / "mul_mat_q4_K_bcast.npz")
    W = g["W"].reshape(1, 2, -1, g["W"].shape[-1])
    got = ctx.mul_mat(int(g["type"]), cu(W), cu(g["X"])).cpu().numpy()
    assert oracle.nmse(g["out"], got) <= NMSE_TIGHT


@pytest.mark.parametrize("f", sorted(glob.glob(str(GOLD / "mul_mat_id_*.npz"))))
def test_mul_mat_id_golden(ctx, oracle, f):
    g = np.load(f)
    got = ctx.mul_mat_id(int(g["type"]), cu(g["W"]), cu(g["B"]), cu(g["ids"])).cpu().numpy()
    assert oracle.nmse(g["out"], got) <= NMSE_TIGHT


# ---------------------------------------------------------------- K3 against the oracle, seeded

SHAPES = [  # (M, K): K/256 = 1, 2, 3 (L=4, one idle lane group), 16 (two rows per warp), 24 (idle lanes), 56 (two slices)
    (67, 256), (130, 512), (33, 768), (96, 4096), (40, 6144), (24, 14336),
]


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
@pytest.mark.parametrize("M,K", SHAPES)
@pytest.mark.parametrize("N", [1, 2, 5, 8])
def test_mul_mat_vs_oracle(ctx, oracle, t, M, K, N):
    if K >= 4096 and N in (2, 5):
        pytest.skip("covered by the smaller K")
    rng = np.random.default_rng(M * 131 + K + N + t)
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, W, X), got) <= NMSE_TIGHT


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
@pytest.mark.parametrize("N", [9, 17, 64])
def test_mul_mat_more_than_8_columns(ctx, oracle, t, N, mmq_bound):
    rng = np.random.default_rng(N + t)
    M, K = 100, 1024
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, W, X), got) <= mmq_bound


@pytest.mark.parametrize("t,K", [(Q4_0, 32), (Q4_0, 96), (Q8_0, 160), (Q8_0, 1056), (Q4_0, 4128)])
def test_generic_kernel_for_k_not_multiple_of_256(ctx, oracle, t, K):
    rng = np.random.default_rng(K)
    M, N = 37, 3
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, W, X), got) <= NMSE_TIGHT


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
def test_fast_and_generic_kernels_agree(ctx, pkg, oracle, t):
    """two independent CUDA implementations of the same integer arithmetic"""
    rng = np.random.default_rng(3 + t)
    M, K, N = 200, 2048, 4
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    a = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    ctx.force_path(pkg.PATH_SIMPLE)
    try:
        b = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    finally:
        ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(a, b) <= NMSE_TIGHT
    assert oracle.nmse(oracle.mul_mat(t, W, X), b) <= NMSE_TIGHT


def test_empty_and_degenerate(ctx, pkg, oracle):
    import torch
    W = quantize_weights(oracle, Q4_0, np.ones((4, 256), dtype=np.float32))
    X = torch.zeros((0, 256), dtype=torch.float32, device="cuda")
    out = ctx.mul_mat(Q4_0, cu(W), X)
    assert out.shape == (0, 4)
    # all-zero activations and all-zero weights give exact zeros
    Z = quantize_weights(oracle, Q4_0, np.zeros((4, 256), dtype=np.float32))
    assert float(ctx.mul_mat(Q4_0, cu(Z), cu(np.ones((2, 256), dtype=np.float32))).abs().max()) == 0.0
    assert float(ctx.mul_mat(Q4_0, cu(W), cu(np.zeros((2, 256), dtype=np.float32))).abs().max()) == 0.0


def test_strided_weight_view(ctx, pkg, oracle):
    """src0 as a view: rows taken out of a wider tensor (nb01 > row size), as the reference's k_v cases do"""
    rng = np.random.default_rng(77)
    M, K, KV, N = 20, 512, 1024, 2
    Wfull = quantize_weights(oracle, Q8_0, rng.uniform(-1, 1, (M, KV)).astype(np.float32))   # rows of KV
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    rb, rbv = pkg.row_size(Q8_0, K), pkg.row_size(Q8_0, KV)
    d = pkg.MulMatDesc.contiguous(Q8_0, M, K, N)
    d.nb01, d.nb02, d.nb03 = rbv, rbv * M, rbv * M
    got = ctx.mul_mat(Q8_0, cu(Wfull), cu(X), desc=d).cpu().numpy()[0, 0]
    ref = oracle.mul_mat(Q8_0, np.ascontiguousarray(Wfull[:, :rb]), X)
    assert oracle.nmse(ref, got) <= NMSE_TIGHT


@pytest.mark.parametrize("t", [Q4_0, Q4_K])
def test_shared_src1_fusion(ctx, oracle, t):
    rng = np.random.default_rng(9)
    K, N = 1024, 2
    Ws = [quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32)) for M in (64, 16, 16, 200, 8)]
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    outs = ctx.mul_mat_shared(t, [cu(W) for W in Ws], cu(X))
    for W, o in zip(Ws, outs):
        assert oracle.nmse(oracle.mul_mat(t, W, X), o.cpu().numpy()) <= NMSE_TIGHT


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
@pytest.mark.parametrize("N", [40, 300])
def test_shared_src1_prefill_group(pkg, oracle, t, N):
    """more than 8 columns and several matrices that read the same src1 (q/k/v, gate/up) with the weight cache on: the first call
    expands every matrix (per-matrix launches, one shared activation operand), later calls run ONE grouped launch of the kind::f16
    GEMM (tiles of all matrices in one stream-K schedule). Both must match the per-matrix results bit for bit and the oracle."""
    rng = np.random.default_rng(90 + t + N)
    K = 1024
    Ws = [quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32)) for M in (640, 130, 128, 257)]
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    Wd, Xd = [cu(W) for W in Ws], cu(X)
    c = pkg.Context(0)
    try:
        c.weight_cache(True)
        single = [c.mul_mat(t, W, Xd).cpu().numpy() for W in Wd]
        first = [o.cpu().numpy() for o in c.mul_mat_shared(t, Wd, Xd)]
        grouped = [o.cpu().numpy() for o in c.mul_mat_shared(t, Wd, Xd)]
        again = [o.cpu().numpy() for o in c.mul_mat_shared(t, Wd[:2], Xd)]
        for i, W in enumerate(Ws):
            assert oracle.nmse(oracle.mul_mat(t, W, X), grouped[i]) <= 1e-6
            assert np.allclose(first[i], single[i], rtol=0, atol=0) or oracle.nmse(single[i], first[i]) <= 1e-12
            assert oracle.nmse(single[i], grouped[i]) <= 1e-12   # the K split of a tile may differ between schedules: fp32 order only
        assert oracle.nmse(single[1], again[1]) <= 1e-12
    finally:
        c.weight_cache_clear()
        c.close()


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
def test_decode_weight_prefetch_changes_nothing(pkg, oracle, t):
    """b200_ctx_set_prefetch: a repeating chain of mat-vecs (what a decode step is) with the next kernel's weights pulled into
    L2 by the current one. Same bits as without it; a chain that changes under the predictor (wrong predictions) too."""
    rng = np.random.default_rng(41 + t)
    shapes = [(96, 4096), (200, 1024), (64, 14336), (33, 512)]
    Ws = [cu(quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))) for (M, K) in shapes]
    Xs = [cu(rng.uniform(-1, 1, (1, K)).astype(np.float32)) for (_, K) in shapes]
    c = pkg.Context(0)
    try:
        base = [c.mul_mat(t, W, X).cpu().numpy() for W, X in zip(Ws, Xs)]
        c.prefetch(128)
        for order in ([0, 1, 2, 3], [0, 1, 2, 3], [3, 1, 0, 2], [0, 0, 2, 2], [0, 1, 2, 3]):
            for i in order:
                got = c.mul_mat(t, Ws[i], Xs[i]).cpu().numpy()
                assert np.array_equal(got, base[i])
        fused = c.mul_mat_shared(t, [Ws[1], Ws[1]], Xs[1])          # a fused launch as predecessor and successor
        again = c.mul_mat_shared(t, [Ws[1], Ws[1]], Xs[1])
        assert np.array_equal(fused[0].cpu().numpy(), base[1]) and np.array_equal(again[1].cpu().numpy(), base[1])
        c.prefetch(0)
        assert np.array_equal(c.mul_mat(t, Ws[2], Xs[2]).cpu().numpy(), base[2])
    finally:
        c.close()
    assert oracle.nmse(oracle.mul_mat(t, Ws[0].cpu().numpy(), Xs[0].cpu().numpy()), base[0]) <= NMSE_TIGHT


# ---------------------------------------------------------------- K5/K6

@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
@pytest.mark.parametrize("n_as,n_used,bcast,n_tok", [(4, 1, 1, 1), (8, 2, 1, 1), (8, 4, 0, 5), (4, 2, 1, 17), (8, 2, 0, 32)])
def test_mul_mat_id_vs_oracle(ctx, oracle, t, n_as, n_used, bcast, n_tok):
    """the shapes of the reference's own cases (tests/test-backend-ops.cpp:7394-7407: m=512, k=256), scaled down in m"""
    rng = np.random.default_rng(n_as * 7 + n_used + n_tok + t)
    M, K = 96, 256
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1 if bcast else n_used, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    got = ctx.mul_mat_id(t, cu(W), cu(B), cu(ids)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat_id(t, W, B, ids), got) <= NMSE_TIGHT


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
@pytest.mark.parametrize("n_as,n_used,bcast,n_tok", [(4, 2, 1, 129), (8, 4, 0, 129), (4, 4, 1, 64), (8, 2, 1, 512), (8, 3, 0, 200)])
def test_mul_mat_id_prefill_batches_per_expert_gemm(ctx, oracle, t, n_as, n_used, bcast, n_tok):
    """from 48 pairs per expert on average: sort, gather, one tile GEMM per expert, un-permuting scatter (the reference's ids helper ->
    mul_mat_q windows -> write-back, ggml-cuda/mmq.cu:148-206). The reference's largest case (n = 129, tests/test-backend-ops.cpp:
    7394-7407) and a 512-token one; skewed routing (one hot expert, one unused); bound = the prefill GEMM's (fp16 operands)."""
    rng = np.random.default_rng(n_as * 11 + n_used + n_tok + t)
    M, K = 160, 512
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1 if bcast else n_used, K)).astype(np.float32)
    n_pool = n_as - 1 if n_used < n_as else n_as                 # the last expert is never chosen (when the routing leaves a choice)
    pool = np.arange(n_pool)
    wts = np.ones(n_pool); wts[0] = 6.0; wts /= wts.sum()        # expert 0 is hot
    ids = np.stack([rng.choice(pool, size=n_used, replace=False, p=wts) for _ in range(n_tok)]).astype(np.int32)
    l0 = ctx.launches
    got = ctx.mul_mat_id(t, cu(W), cu(B), cu(ids)).cpu().numpy()
    assert ctx.launches - l0 >= 4                                # sort, gather, >= 1 GEMM (+ its operand conversion), scatter
    assert oracle.nmse(oracle.mul_mat_id(t, W, B, ids), got) <= NMSE_F16
    again = ctx.mul_mat_id(t, cu(W), cu(B), cu(ids)).cpu().numpy()
    assert np.array_equal(got, again)


# ---------------------------------------------------------------- q5_K (second half: Q5_K_M / Q5_K_S files)

@pytest.mark.parametrize("M,K,N", [(96, 256, 1), (200, 1024, 3), (130, 4096, 1), (64, 4096, 2), (40, 4096, 8), (33, 14336, 1), (257, 1280, 5)])
def test_q5_K_decode(ctx, pkg, oracle, M, K, N):
    """176-byte blocks (q4_K + 32 bytes of fifth bits) are 16-byte multiples: the TMA-staged kernels take any K % 256 == 0; 1-2 columns run
    the dp4a kernel (column c of a 2-column call is bit-identical to the 1-column call), 3-8 the tensor-core mat-vec (same integer sums,
    another fp32 order)"""
    rng = np.random.default_rng(M + K + N + 5)
    W = quantize_weights(oracle, Q5_K, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.standard_normal((N, K)) * rng.choice([0.05, 1.0, 20.0], (N, 1))).astype(np.float32)
    got = ctx.mul_mat(Q5_K, cu(W), cu(X)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(Q5_K, W, X), got) <= NMSE_TIGHT
    if N > 1:
        one = ctx.mul_mat(Q5_K, cu(W), cu(X[1:2])).cpu().numpy()
        assert np.array_equal(one[0], got[1]) if N <= 2 else oracle.nmse(one[0], got[1]) <= 1e-12
    ctx.force_path(pkg.PATH_SIMPLE)
    gen = ctx.mul_mat(Q5_K, cu(W), cu(X)).cpu().numpy()
    ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(gen, got) <= 1e-12


@pytest.mark.parametrize("N", [9, 40, 300])
def test_q5_K_prefill(ctx, oracle, N):
    """more than 8 columns: weights expanded to fp16 (dequantize_row_q5_K values), the kind::f16 GEMM - q5_K has no kind::i8 kernel"""
    rng = np.random.default_rng(N + 5)
    M, K = 300, 1024
    W = quantize_weights(oracle, Q5_K, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.standard_normal((N, K)) * rng.choice([0.01, 1.0, 30.0], (N, 1))).astype(np.float32)
    for f16 in (True, False):   # the switch does not apply to q5_K
        ctx.mmq_f16(f16)
        got = ctx.mul_mat(Q5_K, cu(W), cu(X)).cpu().numpy()
        assert oracle.nmse(oracle.mul_mat(Q5_K, W, X), got) <= NMSE_F16
    ctx.mmq_f16(True)


def test_q5_K_shared_src1_and_fused_glu(ctx, oracle):
    rng = np.random.default_rng(55)
    K = 2048
    Ws = [quantize_weights(oracle, Q5_K, rng.uniform(-1, 1, (m, K)).astype(np.float32)) for m in (64, 17, 130)]
    X = rng.uniform(-1, 1, (2, K)).astype(np.float32)
    outs = ctx.mul_mat_shared(Q5_K, [cu(w) for w in Ws], cu(X))
    for w, o in zip(Ws, outs):
        assert oracle.nmse(oracle.mul_mat(Q5_K, w, X), o.cpu().numpy()) <= NMSE_TIGHT
    Wg = quantize_weights(oracle, Q5_K, rng.uniform(-1, 1, (130, K)).astype(np.float32))
    got = ctx.mul_mat_fused(Q5_K, cu(Ws[2]), cu(Wg), cu(X[:1]), glu_op=2).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat_glu(Q5_K, Ws[2], Wg, X[:1], glu_op=2), got) <= NMSE_TIGHT


# ---------------------------------------------------------------- q6_K (the Q6_K tensors of a Q4_K_M file)

@pytest.mark.parametrize("M,K,N", [(96, 2048, 1), (200, 2048, 3), (130, 4096, 1), (64, 4096, 2), (40, 4096, 8), (33, 14336, 1), (257, 1280, 2)])
def test_q6_K_decode(ctx, pkg, oracle, M, K, N):
    """210-byte blocks are only 2-byte aligned: the TMA-staged kernel (K % 2048 == 0) realigns a lane's unit with funnel shifts,
    other K go through the generic kernel; both hold the integer sums of vec_dot_q6_K_q8_K (ggml-cpu/quants.c:705-758)"""
    rng = np.random.default_rng(M + K + N)
    W = quantize_weights(oracle, Q6_K, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.standard_normal((N, K)) * rng.choice([0.05, 1.0, 20.0], (N, 1))).astype(np.float32)
    got = ctx.mul_mat(Q6_K, cu(W), cu(X)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(Q6_K, W, X), got) <= NMSE_TIGHT
    if N > 1:   # column c of an N-column call equals the 1-column call (from 3 columns on the tensor-core kernel sums a row's
        one = ctx.mul_mat(Q6_K, cu(W), cu(X[1:2])).cpu().numpy()   # units in another fp32 order: same integer sums, NMSE at the fp32 floor)
        if N <= 2 or K % 2048:
            assert np.array_equal(one[0], got[1])
        else:
            assert oracle.nmse(one[0], got[1]) <= 1e-12


@pytest.mark.parametrize("M,K,N", [(16, 2048, 4), (100, 2048, 5), (333, 4096, 8), (48, 14336, 4), (1000, 6144, 7)])
def test_q6_K_tensor_core_matvec(ctx, pkg, oracle, M, K, N):
    """4..8 columns: mma.sync m16n8k16 on signed 6-bit operands, one MMA per 16-weight scale group, units read from 1680-byte lines
    of 8 (2-byte unit phases realigned by funnel shifts); against the oracle, the dp4a kernel and the generic kernel"""
    rng = np.random.default_rng(M * 3 + K + N)
    W = quantize_weights(oracle, Q6_K, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.standard_normal((N, K)) * rng.choice([0.05, 1.0, 20.0], (N, 1))).astype(np.float32)
    ref = oracle.mul_mat(Q6_K, W, X)
    l0 = ctx.launches
    got = ctx.mul_mat(Q6_K, cu(W), cu(X)).cpu().numpy()
    assert ctx.launches - l0 == 2, "quantizer + one tensor-core mat-vec launch"
    assert oracle.nmse(ref, got) <= NMSE_TIGHT
    ctx.force_path(pkg.PATH_SIMPLE)
    gen = ctx.mul_mat(Q6_K, cu(W), cu(X)).cpu().numpy()
    ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(gen, got) <= 1e-12
    two = ctx.mul_mat(Q6_K, cu(W), cu(X[:2])).cpu().numpy()   # dp4a kernel
    assert oracle.nmse(two, got[:2]) <= 1e-12


def test_q6_K_grouped_mul_mat_id(ctx, oracle):
    """enough pairs per expert: pairs sorted by expert, every expert multiplies its tokens 8 at a time on the tensor-core kernel (4-D map)"""
    rng = np.random.default_rng(67)
    n_as, M, K, n_used, n_tok = 4, 80, 2048, 2, 37
    W = quantize_weights(oracle, Q6_K, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    got = ctx.mul_mat_id(Q6_K, cu(W), cu(B), cu(ids)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat_id(Q6_K, W, B, ids), got) <= NMSE_TIGHT


@pytest.mark.parametrize("N", [9, 40, 300])
def test_q6_K_prefill(ctx, oracle, N):
    """more than 8 columns: weights expanded to fp16 (dequantize_row_q6_K values), the kind::f16 GEMM - q6_K has no kind::i8 kernel"""
    rng = np.random.default_rng(N)
    M, K = 300, 1024
    W = quantize_weights(oracle, Q6_K, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.standard_normal((N, K)) * rng.choice([0.01, 1.0, 30.0], (N, 1))).astype(np.float32)
    for f16 in (True, False):   # the switch does not apply to q6_K
        ctx.mmq_f16(f16)
        got = ctx.mul_mat(Q6_K, cu(W), cu(X)).cpu().numpy()
        assert oracle.nmse(oracle.mul_mat(Q6_K, W, X), got) <= NMSE_F16
    ctx.mmq_f16(True)


def test_q6_K_mul_mat_id_and_shared_src1(ctx, oracle):
    rng = np.random.default_rng(66)
    n_as, M, K, n_used, n_tok = 4, 64, 2048, 2, 9
    W = quantize_weights(oracle, Q6_K, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    got = ctx.mul_mat_id(Q6_K, cu(W), cu(B), cu(ids)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat_id(Q6_K, W, B, ids), got) <= NMSE_TIGHT
    Ws = [W[0], W[1][:17], W[2]]
    X = rng.uniform(-1, 1, (2, K)).astype(np.float32)
    outs = ctx.mul_mat_shared(Q6_K, [cu(w) for w in Ws], cu(X))
    for w, o in zip(Ws, outs):
        assert oracle.nmse(oracle.mul_mat(Q6_K, w, X), o.cpu().numpy()) <= NMSE_TIGHT


def test_mm_ids_sort(ctx):
    rng = np.random.default_rng(4)
    for n_as, n_used, n_tok in ((8, 2, 1), (8, 2, 64), (128, 8, 33), (4, 4, 129)):
        ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
        sp, eb = ctx.mm_ids_sort(cu(ids), n_as)
        sp, eb = sp.cpu().numpy(), eb.cpu().numpy()
        flat = ids.reshape(-1)
        order = np.argsort(flat, kind="stable")
        assert np.array_equal(sp, order.astype(np.int32))
        assert np.array_equal(eb, np.concatenate([[0], np.cumsum(np.bincount(flat, minlength=n_as))]).astype(np.int32))


# ---------------------------------------------------------------- tcgen05 mat-vec from raw q4_K blocks (mvtc.cuh)

@pytest.mark.parametrize("M,K,N", [(300, 1024, 4), (128, 4096, 8), (1000, 2048, 9), (129, 14336, 16), (4096, 4096, 13), (77, 512, 17), (640, 8192, 32)])
def test_tcgen05_matvec_from_raw_blocks(pkg, oracle, monkeypatch, M, K, N):
    """mvtc_q4k_kernel over its whole column range (the default only uses it for 9..16 columns): tail tiles (M % 128 != 0), K cut into
    parts whose partial sums meet in the workspace (small M, long K), 8 / 16 / 32-column instantiations. Same integers as ggml-cpu:
    the bound is the decode kernels' 1e-10. Also twice in a row: the part counters must be back at zero."""
    monkeypatch.setenv("GGML_B200_TC_MIN", "4")
    monkeypatch.setenv("GGML_B200_TC_MAX", "32")
    c = pkg.Context(0)
    try:
        rng = np.random.default_rng(M + K + N)
        W = quantize_weights(oracle, Q4_K, rng.uniform(-1, 1, (M, K)).astype(np.float32))
        X = (rng.standard_normal((N, K)) * rng.choice([0.05, 1.0, 20.0], (N, 1))).astype(np.float32)
        ref = oracle.mul_mat(Q4_K, W, X)
        l0 = c.launches
        a = c.mul_mat(Q4_K, cu(W), cu(X)).cpu().numpy()
        assert c.launches - l0 == 2          # the q8_K tile quantizer + the tcgen05 kernel
        b = c.mul_mat(Q4_K, cu(W), cu(X)).cpu().numpy()
        assert oracle.nmse(ref, a) <= NMSE_TIGHT and np.array_equal(a, b)
    finally:
        c.close()


# ---------------------------------------------------------------- exact mode (the arithmetic behind the free-running greedy streams)

@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
def test_exact_mode_is_bit_identical_to_the_generic_vec_dots(pkg, oracle, t):
    """b200_ctx_set_exact: every fp32 operation in ggml-cpu's generic order (oracle.c restates it and tests/test_oracle.py pins that
    restatement bit for bit to the reference's x64 build): the results must be EQUAL, not close - decode, batched decode, prefill
    widths and MUL_MAT_ID. The real-model greedy streams are tests/test_llama_greedy.py."""
    c = pkg.Context(0)
    try:
        c.exact(True)
        rng = np.random.default_rng(100 + t)
        for (M, K, N) in [(67, 2048, 1), (130, 4096, 5), (33, 2048, 40)]:
            W = quantize_weights(oracle, t, (rng.standard_normal((M, K)) * 0.3).astype(np.float32))
            X = (rng.standard_normal((N, K)) * rng.choice([0.01, 1.0, 30.0], (N, 1))).astype(np.float32)
            got = c.mul_mat(t, cu(W), cu(X)).cpu().numpy()
            ref = oracle.mul_mat(t, W, X)
            assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), (TYPE_NAMES[t], M, K, N, float(np.abs(got - ref).max()))
        n_as, M, K, n_used, n_tok = 4, 48, 2048, 2, 5
        W = quantize_weights(oracle, t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
        B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
        ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
        got = c.mul_mat_id(t, cu(W), cu(B), cu(ids)).cpu().numpy()
        ref = oracle.mul_mat_id(t, W, B, ids)
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), TYPE_NAMES[t]
        # the default kernels form the same integers but sum in another order: close, not equal - which is why this mode exists
        c.exact(False)
        W = quantize_weights(oracle, t, (rng.standard_normal((256, 4096)) * 0.3).astype(np.float32))
        X = rng.standard_normal((1, 4096)).astype(np.float32)
        fast = c.mul_mat(t, cu(W), cu(X)).cpu().numpy()
        assert oracle.nmse(oracle.mul_mat(t, W, X), fast) <= NMSE_TIGHT
    finally:
        c.close()


# ---------------------------------------------------------------- full-size properties (sizes the oracle cannot run in seconds)

@pytest.mark.parametrize("t", [Q4_0, Q4_K, Q8_0])
def test_full_size_properties(ctx, pkg, oracle, t):
    """Llama-3-8B FFN shape 14336x4096: (1) a random sample of rows equals the oracle, (2) column c of a 2-column call is
    bit-identical to the 1-column call on that column (same kernel, same summation order) and the 8-column call (tensor-core
    mat-vec for q4 types) agrees with it to rounding, (3) the generic kernel agrees on the whole output."""
    import torch
    rng = np.random.default_rng(31 + t)
    M, K, N = 14336, 4096, 8
    # quantize a slab with the reference quantizer and tile it (quantizing 58M weights on the host takes too long)
    slab = quantize_weights(oracle, t, rng.uniform(-1, 1, (512, K)).astype(np.float32))
    perm = rng.permutation(M) % 512
    W = np.ascontiguousarray(slab[perm])
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    Wd, Xd = cu(W), cu(X)
    full = ctx.mul_mat(t, Wd, Xd)
    rows = rng.choice(M, 64, replace=False)
    ref = oracle.mul_mat(t, np.ascontiguousarray(W[rows]), X)
    assert oracle.nmse(ref, full[:, rows].cpu().numpy()) <= NMSE_TIGHT
    one = ctx.mul_mat(t, Wd, Xd[3:4].contiguous())
    two = ctx.mul_mat(t, Wd, Xd[2:4].contiguous())
    assert torch.equal(one[0], two[1])
    assert oracle.nmse(one[0].cpu().numpy(), full[3].cpu().numpy()) <= NMSE_TIGHT
    ctx.force_path(pkg.PATH_SIMPLE)
    try:
        gen = ctx.mul_mat(t, Wd, Xd)
    finally:
        ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(gen.cpu().numpy(), full.cpu().numpy()) <= NMSE_TIGHT


# ---------------------------------------------------------------- K2/K4: the tcgen05 int8 tile kernel

MMQ_SHAPES = [  # (M, K, N): full tiles, M / N / K tails (K % 128 != 0 -> TMA zero fill), one-block K, one column
    (128, 128, 128), (200, 256, 64), (130, 4096, 33), (512, 1024, 512), (64, 32, 16), (96, 96, 130), (257, 2048, 1), (300, 14336, 40),
]


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
@pytest.mark.parametrize("M,K,N", MMQ_SHAPES)
def test_mmq_vs_oracle(ctx, pkg, oracle, t, M, K, N, mmq_bound):
    if t == Q4_K and K % 256:
        pytest.skip("q4_K needs K % 256 == 0")
    rng = np.random.default_rng(M + 3 * K + 7 * N + t)
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = (rng.standard_normal((N, K)) * rng.choice([0.01, 1.0, 30.0], (N, 1))).astype(np.float32)
    ctx.force_path(pkg.PATH_MMQ)
    try:
        got = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    finally:
        ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(oracle.mul_mat(t, W, X), got) <= mmq_bound


@pytest.mark.parametrize("t", [Q4_0, Q4_K])
def test_mmq_and_matvec_kernels_agree(ctx, pkg, oracle, t, mmq_bound):
    """tensor-core tiles and dp4a mat-vec form the same integer block sums; only the fp32 summation order differs"""
    rng = np.random.default_rng(17 + t)
    M, K, N = 384, 4096, 8
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    a = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    ctx.force_path(pkg.PATH_MMQ)
    try:
        b = ctx.mul_mat(t, cu(W), cu(X)).cpu().numpy()
    finally:
        ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(a, b) <= mmq_bound


def test_mmq_batched_broadcast(ctx, pkg, oracle, mmq_bound):
    rng = np.random.default_rng(23)
    M, K, N = 70, 512, 20
    W = quantize_weights(oracle, Q8_0, rng.uniform(-1, 1, (2 * M, K)).astype(np.float32)).reshape(1, 2, M, -1)
    X = rng.uniform(-1, 1, (1, 4, N, K)).astype(np.float32)
    ctx.force_path(pkg.PATH_MMQ)
    try:
        got = ctx.mul_mat(Q8_0, cu(W), cu(X)).cpu().numpy()
    finally:
        ctx.force_path(pkg.PATH_AUTO)
    ref = oracle.mul_mat(Q8_0, W, X, ne02=2, ne03=1)
    assert oracle.nmse(ref, got) <= mmq_bound


def test_mmq_weight_cache(ctx, pkg, oracle, mmq_bound):
    """a context that keeps int8 expansions must be told about overwritten weights (what the ggml buffer does on set_tensor)"""
    import torch
    rng = np.random.default_rng(29)
    M, K, N = 256, 1024, 48
    W1 = quantize_weights(oracle, Q4_0, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    W2 = quantize_weights(oracle, Q4_0, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    Wd, Xd = cu(W1), cu(X)
    ctx.weight_cache(True)
    try:
        a = ctx.mul_mat(Q4_0, Wd, Xd).cpu().numpy()
        assert pkg.lib().b200_weight_cache_bytes() > 0
        a2 = ctx.mul_mat(Q4_0, Wd, Xd).cpu().numpy()        # served from the cache
        assert np.array_equal(a, a2)
        Wd.copy_(torch.from_numpy(W2))
        pkg.lib().b200_weight_cache_invalidate(0, Wd.data_ptr(), Wd.numel())
        b = ctx.mul_mat(Q4_0, Wd, Xd).cpu().numpy()
    finally:
        ctx.weight_cache(False)
        ctx.weight_cache_clear()
    assert pkg.lib().b200_weight_cache_bytes() == 0
    assert oracle.nmse(oracle.mul_mat(Q4_0, W1, X), a) <= mmq_bound
    assert oracle.nmse(oracle.mul_mat(Q4_0, W2, X), b) <= mmq_bound


def test_mmq_weight_cache_budget_lru_and_view_invalidation(ctx, pkg, oracle):
    """the cache is bounded (least recently used expansions go first, a matrix that cannot be cached is expanded for the
    one call instead of failing), and a write through a view that starts behind row 0 drops the expansion too"""
    import torch
    L = pkg.lib()
    L.b200_weight_cache_set_budget.argtypes = [__import__("ctypes").c_size_t]
    rng = np.random.default_rng(31)
    M, K, N = 256, 1024, 40
    Ws = [quantize_weights(oracle, Q4_K, rng.uniform(-1, 1, (M, K)).astype(np.float32)) for _ in range(4)]
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    Wd, Xd = [cu(w) for w in Ws], cu(X)
    refs = [oracle.mul_mat(Q4_K, w, X) for w in Ws]
    one = M * K * 2                                   # fp16 expansion of one matrix
    ctx.weight_cache(True)
    try:
        L.b200_weight_cache_set_budget(2 * one + 4096)                      # room for two of the four
        for rnd in range(2):
            for i in range(4):
                assert oracle.nmse(refs[i], ctx.mul_mat(Q4_K, Wd[i], Xd).cpu().numpy()) <= NMSE_F16
                assert 0 < L.b200_weight_cache_bytes() <= 2 * one + 4096
        L.b200_weight_cache_set_budget(one // 2)                            # nothing fits: expanded per call, nothing cached, no error
        ctx.weight_cache_clear()
        assert oracle.nmse(refs[0], ctx.mul_mat(Q4_K, Wd[0], Xd).cpu().numpy()) <= NMSE_F16
        assert L.b200_weight_cache_bytes() == 0
        L.b200_weight_cache_set_budget(64 * one)
        a = ctx.mul_mat(Q4_K, Wd[0], Xd).cpu().numpy()
        assert L.b200_weight_cache_bytes() == one
        # overwrite the LAST rows only and report exactly that byte range (a view of the tensor)
        rb = Ws[1].shape[1]
        Wd[0][M - 8:].copy_(torch.from_numpy(Ws[1][M - 8:]))
        L.b200_weight_cache_invalidate(0, Wd[0].data_ptr() + (M - 8) * rb, 8 * rb)
        assert L.b200_weight_cache_bytes() == 0
        b = ctx.mul_mat(Q4_K, Wd[0], Xd).cpu().numpy()
        mixed = np.concatenate([Ws[0][:M - 8], Ws[1][M - 8:]])
        assert oracle.nmse(oracle.mul_mat(Q4_K, mixed, X), b) <= NMSE_F16 and not np.array_equal(a, b)
    finally:
        L.b200_weight_cache_set_budget(0)
        ctx.weight_cache(False)
        ctx.weight_cache_clear()


@pytest.fixture(params=["f16", "i8"])
def mmq_bound(request, ctx):
    """both prefill arithmetics (b200_ctx_set_mmq_f16): scale-carrying fp16 operands (default) and exact per-block int8"""
    ctx.mmq_f16(request.param == "f16")
    yield NMSE_F16 if request.param == "f16" else NMSE_TIGHT
    ctx.mmq_f16(True)


@pytest.mark.parametrize("t", [Q4_0, Q4_K, Q8_0])
def test_mmq_full_size_properties(ctx, pkg, oracle, t, mmq_bound):
    """Llama-3-8B prefill shape 14336x4096, n=512: sampled rows equal the oracle; columns equal the mat-vec kernels' result"""
    rng = np.random.default_rng(37 + t)
    M, K, N = 14336, 4096, 512
    slab = quantize_weights(oracle, t, rng.uniform(-1, 1, (512, K)).astype(np.float32))
    W = np.ascontiguousarray(slab[rng.permutation(M) % 512])
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    Wd, Xd = cu(W), cu(X)
    full = ctx.mul_mat(t, Wd, Xd)                      # n = 512 -> tile kernel
    rows = rng.choice(M, 48, replace=False)
    cols = rng.choice(N, 24, replace=False)
    ref = oracle.mul_mat(t, np.ascontiguousarray(W[rows]), X[cols])
    assert oracle.nmse(ref, full[cols][:, rows].cpu().numpy()) <= mmq_bound
    ctx.force_path(pkg.PATH_MMVQ)
    try:
        mv = ctx.mul_mat(t, Wd, Xd[100:108].contiguous())
    finally:
        ctx.force_path(pkg.PATH_AUTO)
    assert oracle.nmse(mv.cpu().numpy(), full[100:108].cpu().numpy()) <= mmq_bound
