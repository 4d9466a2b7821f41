"""Generates tests/golden/*.npz from the REFERENCE's own compiled code (oracle/_ref, built from /root/reference by
oracle/Makefile): the reference quantizers (ggml-quants.c) and the ggml-cpu x64 (generic scalar) backend running
one-op graphs. Run in the build container:  python tests/golden/make_golden.py  (first set)  and  ... --q6k  (the q6_K files)
The fixtures are small on purpose; they pin oracle/oracle.c and, on the GPU box, the CUDA path."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle.pyoracle import Q4_0, Q4_K, Q6_K, Q8_0, TYPE_NAMES, RefHarness, RefQuant  # noqa: E402

OUT = Path(__file__).resolve().parent


def activations(rng, rows, k):
    x = rng.uniform(-1, 1, (rows, k)).astype(np.float32)
    x[0, :256] = 0.0                      # an all-zero q8_K super-block / eight zero q8_0 blocks
    x[1, 32:64] = 0.0                     # one zero 32-block
    x[2] *= np.float32(1e-6)              # tiny scales (fp16 subnormal d)
    x[3] *= np.float32(3e4)               # large values (d close to fp16 max for q8_0: amax/127 ~ 236)
    x[4, ::2] = np.float32(0.5)           # many ties on the rounding boundary
    x[4, 1::2] = np.float32(-0.5)
    x[5, 5] = np.float32(-7.25)           # signed max is negative -> q8_K iscale positive
    x[6, :] = np.float32(1.0)             # equal |x| everywhere: first-max tie rule of q8_K
    x[6, 17] = np.float32(-1.0)
    return x


def main():
    rq, h = RefQuant(), RefHarness.get()
    cpu = h.cpu("x64")
    rng = np.random.default_rng(20261104)

    x = activations(rng, 12, 1024)
    np.savez_compressed(OUT / "act_quant.npz", x=x, q8_0=rq.quantize_ref("q8_0", x), q8_1=rq.quantize_ref("q8_1", x),
                        q8_K=rq.quantize_ref("q8_K", x))

    for t, (M, K, N) in ((Q4_0, (48, 1024, 5)), (Q8_0, (48, 1024, 5)), (Q4_K, (48, 1024, 5)), (Q4_0, (33, 96, 3)), (Q8_0, (17, 160, 9))):
        w = rng.uniform(-1, 1, (M, K)).astype(np.float32)
        W = rq.quantize_weights(t, w)
        X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
        out, _ = h.mul_mat(cpu, t, W, X, n_threads=1)
        np.savez_compressed(OUT / f"mul_mat_{TYPE_NAMES[t]}_{M}x{K}x{N}.npz", type=t, W=W, X=X, out=out, w_f32=w if t == Q4_0 else np.zeros(0))

    # batch broadcast: src0 [K, M, 2, 1] against src1 [K, N, 4, 2]
    t, M, K, N = Q4_K, 16, 512, 3
    W = rq.quantize_weights(t, rng.uniform(-1, 1, (2 * M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (2, 4, N, K)).astype(np.float32)
    out, _ = h.mul_mat(cpu, t, W, X, ne02=2, ne03=1, n_threads=1)
    np.savez_compressed(OUT / "mul_mat_q4_K_bcast.npz", type=t, W=W, X=X, out=out, ne02=2, ne03=1)

    for t in (Q4_0, Q8_0, Q4_K):
        n_as, M, K, n_used, n_tok = 4, 32, 512, 2, 5
        W = rq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
        for ne11 in (1, n_used):
            B = rng.uniform(-1, 1, (n_tok, ne11, K)).astype(np.float32)
            ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
            out, _ = h.mul_mat_id(cpu, t, W, B, ids, n_threads=1)
            np.savez_compressed(OUT / f"mul_mat_id_{TYPE_NAMES[t]}_b{ne11}.npz", type=t, W=W, B=B, ids=ids, out=out)
    # weight slabs for bench.py and the full-size tests: 2048 segments of 256 reference-quantized U(-1,1) weights per type
    # (a row of any K is a concatenation of segments, because blocks are independent)
    slabs = {}
    for t in (Q4_0, Q8_0, Q4_K):
        slabs[TYPE_NAMES[t]] = rq.quantize_weights(t, rng.uniform(-1, 1, (2048, 256)).astype(np.float32))
    np.savez(OUT / "weight_slabs.npz", **slabs)
    print("golden fixtures written to", OUT)


def main_q6k():
    """the q6_K fixtures (added after the first set; own seed so that the earlier files stay byte-identical)"""
    rq, h = RefQuant(), RefHarness.get()
    cpu = h.cpu("x64")
    rng = np.random.default_rng(20261105)
    t = Q6_K
    for (M, K, N) in ((48, 1024, 5), (19, 256, 2)):
        W = rq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
        X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
        out, _ = h.mul_mat(cpu, t, W, X, n_threads=1)
        np.savez_compressed(OUT / f"mul_mat_{TYPE_NAMES[t]}_{M}x{K}x{N}.npz", type=t, W=W, X=X, out=out, w_f32=np.zeros(0))
    n_as, M, K, n_used, n_tok = 4, 32, 512, 2, 5
    W = rq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    out, _ = h.mul_mat_id(cpu, t, W, B, ids, n_threads=1)
    np.savez_compressed(OUT / f"mul_mat_id_{TYPE_NAMES[t]}_b1.npz", type=t, W=W, B=B, ids=ids, out=out)
    np.savez(OUT / "weight_slabs_q6_K.npz", q6_K=rq.quantize_weights(t, rng.uniform(-1, 1, (2048, 256)).astype(np.float32)))
    print("q6_K golden fixtures written to", OUT)


def main_q5k():
    """the q5_K fixtures (own seed: the earlier files stay byte-identical)"""
    from oracle.pyoracle import Q5_K
    rq, h = RefQuant(), RefHarness.get()
    cpu = h.cpu("x64")
    rng = np.random.default_rng(20261107)
    t = Q5_K
    for (M, K, N) in ((48, 1024, 5), (19, 256, 2)):
        W = rq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
        X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
        out, _ = h.mul_mat(cpu, t, W, X, n_threads=1)
        np.savez_compressed(OUT / f"mul_mat_{TYPE_NAMES[t]}_{M}x{K}x{N}.npz", type=t, W=W, X=X, out=out, w_f32=np.zeros(0))
    n_as, M, K, n_used, n_tok = 4, 32, 512, 2, 5
    W = rq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    out, _ = h.mul_mat_id(cpu, t, W, B, ids, n_threads=1)
    np.savez_compressed(OUT / f"mul_mat_id_{TYPE_NAMES[t]}_b1.npz", type=t, W=W, B=B, ids=ids, out=out)
    np.savez(OUT / "weight_slabs_q5_K.npz", q5_K=rq.quantize_weights(t, rng.uniform(-1, 1, (2048, 256)).astype(np.float32)))
    print("q5_K golden fixtures written to", OUT)


def main_fusion():
    """K7 fixtures: the fused patterns of test_mul_mat_vec_fusion (tests/test-backend-ops.cpp:5120-5249) run as whole graphs on the
    reference's ggml-cpu (x64 generic build). Own seed: the earlier files stay byte-identical."""
    rq, h = RefQuant(), RefHarness.get()
    cpu = h.cpu("x64")
    rng = np.random.default_rng(20261106)
    cases = {}
    M, K = 96, 512
    for t in (Q4_0, Q8_0, Q4_K, Q6_K):
        Wu = rq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
        Wg = rq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
        X = rng.uniform(-1, 1, (1, K)).astype(np.float32)
        bu = rng.uniform(-1, 1, M).astype(np.float32)
        bg = rng.uniform(-1, 1, M).astype(np.float32)
        tn = TYPE_NAMES[t]
        cases[f"{tn}_Wu"], cases[f"{tn}_Wg"], cases[f"{tn}_X"], cases[f"{tn}_bu"], cases[f"{tn}_bg"] = Wu, Wg, X, bu, bg
        for op in range(6):
            cases[f"{tn}_glu{op}"] = h.mul_mat_glu(cpu, t, Wu, Wg, X, glu_op=op, n_threads=1)
            cases[f"{tn}_glu{op}_bias"] = h.mul_mat_glu(cpu, t, Wu, Wg, X, bu, bg, glu_op=op, n_threads=1)
        cases[f"{tn}_bias_only"] = h.mul_mat_glu(cpu, t, Wu, None, X, bu, None, glu_op=-1, n_threads=1)
    # MUL_MAT_ID / ADD_ID: one token, 4 of 6 experts
    t, n_as, n_used = Q4_K, 6, 4
    Wu = rq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    Wg = rq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (1, 1, K)).astype(np.float32)
    ids = rng.permutation(n_as)[:n_used].astype(np.int32).reshape(1, n_used)
    bu = rng.uniform(-1, 1, (n_as, M)).astype(np.float32)
    bg = rng.uniform(-1, 1, (n_as, M)).astype(np.float32)
    cases.update(id_Wu=Wu, id_Wg=Wg, id_B=B, id_ids=ids, id_bu=bu, id_bg=bg)
    for op in (1, 2, 3):
        cases[f"id_glu{op}"] = h.mul_mat_id_glu(cpu, t, Wu, Wg, B, ids, glu_op=op, n_threads=1)
        cases[f"id_glu{op}_bias"] = h.mul_mat_id_glu(cpu, t, Wu, Wg, B, ids, bu, bg, glu_op=op, n_threads=1)
    cases["id_bias_only"] = h.mul_mat_id_glu(cpu, t, Wu, None, B, ids, bu, None, glu_op=-1, n_threads=1)
    np.savez_compressed(OUT / "fusion_k7.npz", **cases)
    print("K7 golden fixture written to", OUT)


if __name__ == "__main__":
    main_q5k() if "--q5k" in sys.argv else main_fusion() if "--fusion" in sys.argv else main_q6k() if "--q6k" in sys.argv else main()
