"""SURVEY 8(f) rank 3, first slice: the ops between the matmuls of a Llama decode graph (csrc/ops.cu) through the C-ABI, against
float64 restatements of ggml-cpu's formulas (ggml/src/ggml-cpu/ops.cpp: rms_norm :3646-3695, rope :5525-5780, set_rows :4828-4876,
get_rows, dup, flash_attn_ext :7994-8218) and, for the quantized GET_ROWS, bit for bit against the reference's dequantize_row_<type>.
The reference's own cases for the same ops run through test-backend-ops in test_gpu_backend.py; the whole graph runs through
libllama in test_llama_greedy.py."""
import math

import numpy as np
import pytest

from oracle.pyoracle import Q4_0, Q4_K, Q5_K, Q6_K, Q8_0

pytestmark = pytest.mark.gpu


def _nmse(ref, got):
    ref = np.asarray(ref, dtype=np.float64)
    got = np.asarray(got, dtype=np.float64)
    return float(((ref - got) ** 2).sum() / max((ref ** 2).sum(), 1e-30))


@pytest.mark.parametrize("shape,eps", [((1, 1, 1, 4096), 1e-5), ((2, 3, 7, 64), 1e-6), ((1, 1, 40, 8192), 1e-5), ((1, 1, 5, 10), 0.0), ((1, 2, 3, 1), 1e-6)])
def test_rms_norm(ctx, shape, eps):
    import torch
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(shape, device="cuda", generator=g) * 3 + 0.5
    got = ctx.rms_norm(x, eps)
    x64 = x.double()
    ref = x64 / torch.sqrt((x64 * x64).mean(-1, keepdim=True) + eps)
    assert _nmse(ref.cpu().numpy(), got.cpu().numpy()) < 1e-12
    # a strided view (rows of a wider buffer) into a strided destination
    wide = torch.randn(shape[:-1] + (shape[-1] * 2,), device="cuda", generator=g)
    view = wide[..., : shape[-1]]
    out = torch.zeros_like(wide)
    ctx.rms_norm(view, eps, out=out[..., shape[-1]:])
    v64 = view.double()
    ref = v64 / torch.sqrt((v64 * v64).mean(-1, keepdim=True) + eps)
    assert _nmse(ref.cpu().numpy(), out[..., shape[-1]:].cpu().numpy()) < 1e-12
    assert float(out[..., : shape[-1]].abs().max()) == 0.0


@pytest.mark.parametrize("shape", [(1, 1, 1, 4096), (1, 1, 7, 4096), (2, 3, 5, 100), (1, 1, 3, 20000)])
def test_rms_norm_mul_is_the_two_nodes(ctx, shape):
    """build_norm's RMS_NORM + MUL(norm weight) as one launch: bit-identical to running the two ops (vector and scalar paths)"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(shape, device="cuda", generator=g) * 2
    w = torch.randn(shape[-1], device="cuda", generator=g)
    fused = ctx.rms_norm_mul(x, w, 1e-5)
    two = ctx.mul(ctx.rms_norm(x, 1e-5), w.reshape(1, 1, 1, -1))
    assert torch.equal(fused, two)
    x64 = x.double()
    ref = x64 / torch.sqrt((x64 * x64).mean(-1, keepdim=True) + 1e-5) * w.double()
    assert _nmse(ref.cpu().numpy(), fused.cpu().numpy()) < 1e-12


@pytest.mark.parametrize("t", [Q4_K, Q4_0, Q8_0, Q6_K])
@pytest.mark.parametrize("K,N", [(4096, 1), (2048, 1), (4096, 3), (16384, 1)])
def test_norm_folded_into_the_mat_vec_is_the_nodes(pkg, ctx, refq, t, K, N):
    """attn_norm -> q/k/v and ffn_norm -> gate/up with the RMS_NORM + MUL inside the mat-vec kernel's quantizer prologue (one column,
    formats of the lean / GLU kernels) or as its own launch in front (everything else): bit-identical to norm, mul, MUL_MAT(s), GLU
    node by node - the kernels share one summation order (common.cuh unit_sumsq)"""
    import torch
    rng = np.random.default_rng(K + N + t)
    Ms = [512, 128, 128]
    Ws = [torch.from_numpy(refq.quantize_weights(t, rng.normal(0, 1, (M, K)).astype(np.float32) / 8)).cuda() for M in Ms]
    x = torch.from_numpy(rng.normal(0, 2, (N, K)).astype(np.float32)).cuda()
    w = torch.from_numpy(rng.normal(1, 0.2, K).astype(np.float32)).cuda()
    xn = ctx.rms_norm_mul(x.reshape(1, 1, N, K), w, 1e-5).reshape(N, K)
    n0 = ctx.launches
    fused = ctx.mul_mat_shared(t, Ws, x, norm_w=w, norm_eps=1e-5)
    n_fused = ctx.launches - n0
    plain = ctx.mul_mat_shared(t, Ws, xn)
    for a, b in zip(fused, plain):
        assert torch.equal(a, b)
    if N == 1 and t in (Q4_K, Q4_0, Q8_0) and K <= 8192:     # a quantizing warp holds two units: rows up to 8192 elements
        assert n_fused == 1, n_fused          # the norm really ran inside the mat-vec launch
    if N == 1:
        g = ctx.mul_mat_fused(t, Ws[1], Ws[2], x, glu_op=pkg.GLU_SWIGLU, norm_w=w, norm_eps=1e-5)
        g_ref = ctx.mul_mat_fused(t, Ws[1], Ws[2], xn, glu_op=pkg.GLU_SWIGLU)
        assert torch.equal(g, g_ref)


@pytest.mark.parametrize("cache_dtype", ["float16", "float32"])
@pytest.mark.parametrize("n_tok", [1, 5, 37])
def test_attn_prep_is_the_four_nodes(ctx, n_tok, cache_dtype):
    """ROPE(q), ROPE(k), SET_ROWS(k cache), SET_ROWS(v cache) of one layer in one launch: bit-identical to the four ops"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(8 + n_tok)
    D, nh, nkv, n_ctx = 128, 32, 8, 128
    q = torch.randn((1, n_tok, nh, D), device="cuda", generator=g)
    k = torch.randn((1, n_tok, nkv, D), device="cuda", generator=g)
    v = torch.randn((n_tok, nkv * D), device="cuda", generator=g)
    pos = torch.arange(40, 40 + n_tok, device="cuda", dtype=torch.int32)
    idx = (torch.arange(n_tok, device="cuda", dtype=torch.int64) * 3 + 11) % n_ctx
    dt = getattr(torch, cache_dtype)
    kc1, vc1 = torch.zeros((n_ctx, nkv * D), device="cuda", dtype=dt), torch.zeros((n_ctx, nkv * D), device="cuda", dtype=dt)
    kc2, vc2 = torch.zeros_like(kc1), torch.zeros_like(vc1)
    kw = dict(freq_base=500000.0, n_ctx_orig=8192)
    qo, ko = ctx.attn_prep(q, k, pos, kc1, idx, v, vc1, idx, D, 0, **kw)
    q_ref = ctx.rope(q, pos, D, 0, **kw)
    k_ref = ctx.rope(k, pos, D, 0, **kw)
    ctx.set_rows(kc2, k_ref.reshape(n_tok, nkv * D), idx)
    ctx.set_rows(vc2, v, idx)
    assert torch.equal(qo, q_ref) and torch.equal(ko, k_ref)
    assert torch.equal(kc1, kc2) and torch.equal(vc1, vc2)
    assert float(kc1.float().abs().sum()) > 0


@pytest.mark.parametrize("t", [Q4_K, Q4_0, Q8_0, Q6_K])
@pytest.mark.parametrize("K,D,nh,nkv,N,mode,norm", [(4096, 128, 32, 8, 1, 0, True), (4096, 128, 32, 8, 1, 0, False), (2048, 64, 8, 2, 1, 0, True),
                                                    (4096, 128, 32, 8, 1, 2, True), (4096, 128, 32, 8, 3, 0, True), (8192, 128, 8, 8, 1, 0, True)])
def test_qkv_rope_is_the_nodes(pkg, ctx, refq, t, K, D, nh, nkv, N, mode, norm):
    """[RMS_NORM + MUL] + the three attention projections + ROPE(q), ROPE(k) + the KV-cache rows through b200_qkv_rope: ONE launch for one
    token on the lean mat-vec kernel (normal-mode rope: a row group is whole rope pairs), projections + attn_prep otherwise - the rotated
    q / k and both caches bit-identical to norm, MUL_MATs, ROPE, ROPE, SET_ROWS, SET_ROWS node by node"""
    import torch
    rng = np.random.default_rng(K + D + N + t + mode)
    Ws = [torch.from_numpy(refq.quantize_weights(t, rng.normal(0, 1, (M, K)).astype(np.float32) / 8)).cuda() for M in (nh * D, nkv * D, nkv * D)]
    x = torch.from_numpy(rng.normal(0, 2, (N, K)).astype(np.float32)).cuda()
    w = torch.from_numpy(rng.normal(1, 0.2, K).astype(np.float32)).cuda()
    n_ctx = 64
    pos = torch.arange(17, 17 + N, device="cuda", dtype=torch.int32)
    idx = torch.arange(20, 20 + N, device="cuda", dtype=torch.int64)
    kc1 = (torch.randn((n_ctx, nkv * D), device="cuda") * 0.5).half()
    vc1 = (torch.randn((n_ctx, nkv * D), device="cuda") * 0.5).half()
    kc2, vc2 = kc1.clone(), vc1.clone()
    kw = dict(freq_base=500000.0, n_ctx_orig=8192)
    xn = ctx.rms_norm_mul(x.reshape(1, 1, N, K), w, 1e-5).reshape(N, K) if norm else x
    q, k, v = ctx.mul_mat_shared(t, Ws, xn)
    qo_ref, ko_ref = ctx.attn_prep(q.view(1, N, nh, D), k.view(1, N, nkv, D), pos, kc1, idx, v, vc1, idx, D, mode, **kw)
    n0 = ctx.launches
    qo, ko = ctx.qkv_rope(t, Ws, x, pos, kc2, idx, vc2, idx, nh, nkv, D, mode, norm_w=w if norm else None, norm_eps=1e-5 if norm else None, **kw)
    n_launch = ctx.launches - n0
    assert torch.equal(kc2, kc1) and torch.equal(vc2, vc1)
    assert torch.equal(qo, qo_ref) and torch.equal(ko, ko_ref)
    if N == 1 and mode == 0 and t in (Q4_K, Q4_0, Q8_0) and K <= 4096:
        assert n_launch == 1, n_launch      # everything really happened in the mat-vec launch


@pytest.mark.parametrize("mode", [0, 2])
@pytest.mark.parametrize("D,nh,nkv,n_kv,slot", [(128, 32, 8, 256, 70), (128, 32, 8, 1024, 1023), (64, 8, 2, 96, 0), (256, 4, 4, 640, 300), (128, 8, 1, 2048, 1500)])
def test_attn_decode_is_prep_then_flash_attn(ctx, mode, D, nh, nkv, n_kv, slot):
    """one token: ROPE(q), ROPE(k), the two KV-cache writes and FLASH_ATTN_EXT as ONE launch (the kernel rotates q and the new K row itself
    and uses the new rows instead of the cache slot being written) - every output bit-identical to b200_attn_prep + b200_flash_attn_ext"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(D + nh + n_kv + mode)
    n_ctx = n_kv + 64
    q = torch.randn((1, 1, nh, D), device="cuda", generator=g)
    k = torch.randn((1, 1, nkv, D), device="cuda", generator=g)
    v = torch.randn((1, nkv * D), device="cuda", generator=g)
    kc = (torch.randn((n_ctx, nkv * D), device="cuda", generator=g) * 0.5).half()
    vc = (torch.randn((n_ctx, nkv * D), device="cuda", generator=g) * 0.5).half()
    pos = torch.tensor([slot + 5], device="cuda", dtype=torch.int32)
    idx = torch.tensor([slot], device="cuda", dtype=torch.int64)
    mask = torch.zeros((1, 1, 64, n_kv), device="cuda")
    mask[0, 0, :, slot + 1:] = -math.inf      # cells behind the token's own are free (or another sequence's)
    mask[0, 0, :, 2:5] = -math.inf
    mask = mask.half()
    scale = 1.0 / math.sqrt(D)
    kw = dict(freq_base=500000.0, n_ctx_orig=8192)
    kc1, vc1, kc2, vc2 = kc.clone(), vc.clone(), kc.clone(), vc.clone()
    qo_ref, ko_ref = ctx.attn_prep(q, k, pos, kc1, idx, v, vc1, idx, D, mode, **kw)
    kview = kc1[:n_kv].view(1, n_kv, nkv, D).permute(0, 2, 1, 3)
    vview = vc1[:n_kv].view(1, n_kv, nkv, D).permute(0, 2, 1, 3)
    ref = ctx.flash_attn_ext(qo_ref.permute(0, 2, 1, 3), kview, vview, mask, None, scale)
    # the attention behind the cache write, told which row is in flight (with GGML_B200_FA_EARLY=1 it reads the rest ahead of the stream dependency)
    kc3, vc3 = kc.clone(), vc.clone()
    qo3, _ = ctx.attn_prep(q, k, pos, kc3, idx, v, vc3, idx, D, mode, **kw)
    early = ctx.flash_attn_ext(qo3.permute(0, 2, 1, 3), kc3[:n_kv].view(1, n_kv, nkv, D).permute(0, 2, 1, 3), vc3[:n_kv].view(1, n_kv, nkv, D).permute(0, 2, 1, 3),
                               mask, None, scale, written_idx=idx)
    assert torch.equal(early, ref)
    got, qo, ko = ctx.attn_decode(q, k, pos, kc2, idx, v, vc2, idx, n_kv, mask, D, mode, scale=scale, **kw)
    assert torch.equal(qo, qo_ref) and torch.equal(ko, ko_ref)
    assert torch.equal(kc2, kc1) and torch.equal(vc2, vc1)
    assert torch.equal(got, ref)
    assert float(got.abs().sum()) > 0


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
def test_get_rows_quantized_is_dequantize_row(pkg, ctx, refq, t):
    """token-embedding lookup from a quantized tok_embd: bit-identical to dequantize_row_<type> (ggml-quants.c) of the chosen rows"""
    import torch
    rng = np.random.default_rng(t)
    K, M = 1024, 300
    W = refq.quantize_weights(t, rng.normal(0, 1, (M, K)).astype(np.float32))
    ids = rng.integers(0, M, 51).astype(np.int32)
    Wd = torch.from_numpy(W).cuda()
    got = ctx.get_rows(pkg.Tensor.quantized(Wd, t, K, M), torch.from_numpy(ids).cuda())
    ref = refq.dequantize(t, W, K)[ids]
    assert np.array_equal(got.cpu().numpy().reshape(-1, K).view(np.uint32), ref.view(np.uint32))


def test_get_rows_f32_f16_batched(ctx):
    import torch
    g = torch.Generator(device="cuda").manual_seed(2)
    for dt in (torch.float32, torch.float16):
        src = torch.randn((2, 3, 50, 96), device="cuda", generator=g).to(dt)     # [ne03, ne02, ne01, ne00]
        ids = torch.randint(0, 50, (2, 3, 9), device="cuda", generator=g, dtype=torch.int32)   # [ne12 = ne03, ne11 = ne02, ne10]
        got = ctx.get_rows(src, ids)
        ref = torch.gather(src.float(), 2, ids.long().unsqueeze(-1).expand(-1, -1, -1, 96))
        assert torch.equal(got, ref)


def _rope_ref(x, pos, n_dims, mode, ff, n_ctx_orig, freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow):
    """ggml_compute_forward_rope_f32 + rope_yarn + ggml_rope_yarn_corr_dims in float64; x [nb, nt, nh, D]"""
    x = x.astype(np.float64)
    out = x.copy()
    theta_scale = freq_base ** (-2.0 / n_dims)

    def corr_dim(n_rot):
        return n_dims * math.log(n_ctx_orig / (n_rot * 2 * math.pi)) / (2 * math.log(freq_base))
    lo = max(0.0, math.floor(corr_dim(beta_fast))) if n_ctx_orig else 0.0
    hi = min(n_dims - 1.0, math.ceil(corr_dim(beta_slow))) if n_ctx_orig else 0.0
    for it in range(x.shape[1]):
        tb = np.float32(pos[it])   # ggml_rope_cache_init: theta *= theta_scale per pair, in fp32
        for j in range(n_dims // 2):
            extrap = float(tb / (np.float32(ff[j]) if ff is not None else np.float32(1.0)))
            tb = np.float32(tb * np.float32(theta_scale))
            interp = freq_scale * extrap
            theta, m = interp, attn_factor
            if ext_factor != 0.0:
                y = (j - lo) / max(0.001, hi - lo)
                mix = (1 - min(1.0, max(0.0, y))) * ext_factor
                theta = interp * (1 - mix) + extrap * mix
                m *= 1 + 0.1 * math.log(1 / freq_scale)
            c, s = math.cos(theta) * m, math.sin(theta) * m
            ia, ib = (j, j + n_dims // 2) if mode == 2 else (2 * j, 2 * j + 1)
            a, b = x[:, it, :, ia], x[:, it, :, ib]
            out[:, it, :, ia] = a * c - b * s
            out[:, it, :, ib] = a * s + b * c
    return out


@pytest.mark.parametrize("mode", [0, 2])
@pytest.mark.parametrize("D,n_dims,use_ff,yarn", [(128, 128, False, False), (128, 128, True, False), (64, 32, False, False), (128, 128, False, True), (80, 80, True, True)])
def test_rope(ctx, mode, D, n_dims, use_ff, yarn):
    import torch
    rng = np.random.default_rng(D + n_dims + mode)
    x = rng.normal(0, 1, (2, 5, 4, D)).astype(np.float32)
    pos = np.array([0, 1, 7, 300, 4095], dtype=np.int32)
    ff = rng.uniform(0.9, 1.1, n_dims // 2).astype(np.float32) if use_ff else None
    kw = dict(n_ctx_orig=8192, freq_base=500000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0, beta_fast=32.0, beta_slow=1.0)
    if yarn:
        kw.update(freq_scale=0.25, ext_factor=1.0, attn_factor=1.2, n_ctx_orig=4096, freq_base=10000.0)
    got = ctx.rope(torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), n_dims, mode, None if ff is None else torch.from_numpy(ff).cuda(), **kw)
    ref = _rope_ref(x, pos, n_dims, mode, ff, **kw)
    # the angle is formed with the reference's fp32 products; what is left is sinf / cosf and the fp32 rotation
    assert _nmse(ref, got.cpu().numpy()) < 1e-10
    if n_dims < D:
        assert np.array_equal(got.cpu().numpy()[..., n_dims:], x[..., n_dims:])


@pytest.mark.parametrize("dst_dtype", ["float16", "float32"])
@pytest.mark.parametrize("idx_dtype", ["int64", "int32"])
def test_set_rows_kv_cache_write(ctx, dst_dtype, idx_dtype):
    """llama_kv_cache::cpy_k (src/llama-kv-cache.cpp): n_tokens F32 rows of n_embd_k_gqa into the F16 cache at the slot indices"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(3)
    n_tok, n_embd, n_ctx = 7, 1024, 64
    src = torch.randn((1, 1, n_tok, n_embd), device="cuda", generator=g)
    idx = torch.tensor([5, 6, 7, 40, 0, 63, 12], device="cuda", dtype=getattr(torch, idx_dtype))
    dst = torch.zeros((1, 1, n_ctx, n_embd), device="cuda", dtype=getattr(torch, dst_dtype))
    ctx.set_rows(dst, src, idx)
    ref = torch.zeros_like(dst)
    ref[0, 0, idx.long()] = src[0, 0].to(dst.dtype)
    assert torch.equal(dst, ref)
    # broadcast over ne2 (ne11 == 1): every plane of src goes through the same index row
    src = torch.randn((1, 3, n_tok, 96), device="cuda", generator=g)
    dst = torch.zeros((1, 3, n_ctx, 96), device="cuda", dtype=getattr(torch, dst_dtype))
    ctx.set_rows(dst, src, idx.reshape(1, 1, n_tok))
    ref = torch.zeros_like(dst)
    ref[0, :, idx.long()] = src[0].to(dst.dtype)
    assert torch.equal(dst, ref)


def test_cpy_cont_casts_and_strides(ctx):
    import torch
    g = torch.Generator(device="cuda").manual_seed(4)
    a = torch.randn((2, 3, 5, 32), device="cuda", generator=g)
    # F32 -> F16 cast (the KQ mask of build_attn_mha with flash attention, src/llama-graph.cpp:1343-1345)
    h = torch.empty(a.shape, device="cuda", dtype=torch.float16)
    ctx.cpy(a, h)
    assert torch.equal(h, a.half())
    # CONT of a permuted view (q / k / v head permutes)
    p = a.permute(0, 2, 1, 3)
    c = torch.empty(p.shape, device="cuda")
    ctx.cpy(p, c)
    assert torch.equal(c, p.contiguous())
    # reshaping copy: same element count, other shape, into a strided destination
    wide = torch.zeros((30, 64), device="cuda", dtype=torch.float16)
    ctx.cpy(p, wide[:, :32])
    assert torch.equal(wide[:, :32], p.contiguous().reshape(30, 32).half())
    assert float(wide[:, 32:].abs().max()) == 0.0
    back = torch.empty((30, 32), device="cuda")
    ctx.cpy(wide[:, :32], back)
    assert torch.equal(back, wide[:, :32].float())


def _fa_ref(q, k, v, mask, sinks, scale, max_bias, softcap):
    import torch
    nb_, nh, nq, D = q.shape
    nkv_h = k.shape[1]
    q64 = q.half().double()      # ggml-cpu converts q to fp16 for the F16-K dot product (ops.cpp:8109-8110)
    k64 = k.double().repeat_interleave(nh // nkv_h, dim=1)
    v64 = v.double().repeat_interleave(nh // v.shape[1], dim=1)
    s = torch.einsum("bhqd,bhkd->bhqk", q64, k64) * (scale / softcap if softcap else scale)
    if softcap:
        s = softcap * torch.tanh(s)
    if mask is not None:
        m = mask[:, :, :nq, :].double()
        if max_bias > 0:
            n2 = 1 << int(math.floor(math.log2(nh)))
            m0, m1 = 2.0 ** (-max_bias / n2), 2.0 ** (-max_bias / 2 / n2)
            slope = torch.tensor([m0 ** (h + 1) if h < n2 else m1 ** (2 * (h - n2) + 1) for h in range(nh)], dtype=torch.float64, device=q.device)
            m = m * slope.view(1, nh, 1, 1)
        s = s + m
    if sinks is not None:
        s = torch.cat([s, sinks.double().view(1, nh, 1, 1).expand(nb_, nh, nq, 1)], dim=-1)
    p = torch.softmax(s, dim=-1)
    if sinks is not None:
        p = p[..., :-1]
    return torch.einsum("bhqk,bhkd->bqhd", p, v64)


@pytest.mark.parametrize("D,nh,nkv_h,nq,n_kv,opts", [
    (128, 32, 8, 1, 256, {}),                 # Llama-3-8B decode, first cache page
    (128, 32, 8, 1, 4096, {}),                # ... a long context: KV splits + combine
    (128, 32, 8, 64, 256, {}),                # the 64-token prompt of the greedy test: tensor-core tiles
    (128, 8, 2, 200, 456, {}),                # a ragged prompt batch on top of 256 cached positions (partial row block, partial KV tile)
    (64, 4, 4, 33, 97, {"sinks": True}),
    (128, 4, 1, 17, 130, {"softcap": 10.0}),
    (128, 4, 2, 40, 300, {"max_bias": 8.0}),
    (256, 4, 2, 20, 150, {}),                 # head size 256: the row-block kernel
    (128, 4, 2, 9, 200, {}),                  # 4..15 rows: the row-block kernel
    (64, 4, 2, 3, 113, {"sinks": True}),
    (256, 8, 8, 2, 1024, {"softcap": 10.0}),
    (128, 8, 2, 5, 777, {"max_bias": 8.0}),
    (128, 6, 3, 1, 1100, {"nomask": True, "sinks": True}),
])
def test_flash_attn_ext(ctx, D, nh, nkv_h, nq, n_kv, opts):
    import torch
    g = torch.Generator(device="cuda").manual_seed(D + nq + n_kv)
    q = torch.randn((1, nh, nq, D), device="cuda", generator=g)
    # K / V as views of a bigger cache, as llama_kv_cache::get_k hands them out: [n_kv, n_head_kv, D] permuted to heads-major
    cache_k = torch.randn((1, n_kv + 64, nkv_h, D), device="cuda", generator=g).half()
    cache_v = torch.randn((1, n_kv + 64, nkv_h, D), device="cuda", generator=g).half()
    k = cache_k[:, :n_kv].permute(0, 2, 1, 3)
    v = cache_v[:, :n_kv].permute(0, 2, 1, 3)
    mask = None
    if not opts.get("nomask"):
        nq_pad = (nq + 63) // 64 * 64
        mask = torch.zeros((1, 1, nq_pad, n_kv), device="cuda")
        causal_from = n_kv - nq     # query i sees positions <= causal_from + i
        for i in range(nq):
            mask[0, 0, i, causal_from + i + 1:] = -math.inf
        mask[0, 0, :, 3:9] = -math.inf    # a hole (another sequence's cells)
        if opts.get("max_bias"):
            mask = torch.where(torch.isinf(mask), mask, -torch.rand(mask.shape, device="cuda", generator=g))
        mask = mask.half()
    sinks = torch.randn(nh, device="cuda", generator=g) * 3 if opts.get("sinks") else None
    scale = 1.0 / math.sqrt(D)
    got = ctx.flash_attn_ext(q, k, v, mask, sinks, scale, opts.get("max_bias", 0.0), opts.get("softcap", 0.0))
    ref = _fa_ref(q, k, v, mask, sinks, scale, opts.get("max_bias", 0.0), opts.get("softcap", 0.0))
    assert got.shape == ref.shape
    # decode / row-block kernels: fp32 accumulation of exact fp16 products. Tile kernel (16 rows and more, D <= 128): the softmax weights
    # are rounded to fp16 for the tensor-core P V product (ggml-cpu rounds the whole accumulator to fp16). test-backend-ops' bar is 5e-4.
    assert _nmse(ref.cpu().numpy(), got.cpu().numpy()) < (1e-10 if nq < 16 or D > 128 else 2e-6)


def test_flash_attn_ext_all_masked_row_is_zero(ctx):
    import torch
    q = torch.ones((1, 2, 1, 64), device="cuda")
    k = torch.ones((1, 2, 32, 64), device="cuda", dtype=torch.float16)
    mask = torch.full((1, 1, 64, 32), -math.inf, device="cuda", dtype=torch.float16)
    got = ctx.flash_attn_ext(q, k, k, mask, None, 1.0)
    assert float(got.abs().max()) == 0.0   # S == 0 -> 0 (ops.cpp:8207)


def test_unsupported_layouts_are_refused(pkg, ctx):
    import torch
    q = torch.ones((1, 2, 1, 80), device="cuda")
    k = torch.ones((1, 2, 32, 80), device="cuda", dtype=torch.float16)
    with pytest.raises(Exception):
        ctx.flash_attn_ext(q, k, k, None, None, 1.0)     # head size 80: not claimed by supports_op either
    x = torch.ones((1, 1, 4, 64), device="cuda")
    with pytest.raises(Exception):
        ctx.rope(x, torch.zeros(4, device="cuda", dtype=torch.int32), 64, mode=8)   # mrope
