"""GPU drop-in tests (pytest -m gpu): the ggml backend library loaded through the reference's own registry
(ggml_backend_load -> ggml_backend_init) and driven through the reference's public API, plus the reference's
conformance harness tests/test-backend-ops.cpp (compiled into oracle/_ref) pointed at it with GGML_BACKEND_PATH."""
import os
import re
import subprocess

import numpy as np
import pytest

from oracle.pyoracle import Q4_0, Q4_K, Q5_K, Q6_K, Q8_0, REF_DIR, have_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def b200(pkg, harness):
    if not pkg.BACKEND_PATH.exists():
        pytest.fail("libggml-b200.so missing: build it with __graft_entry__.build()")
    return harness.load(pkg.BACKEND_PATH)


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K])
@pytest.mark.parametrize("M,K,N", [(64, 256, 1), (128, 1024, 4), (96, 4096, 1), (50, 512, 33)])
def test_backend_mul_mat_matches_ggml_cpu(harness, refq, oracle, b200, t, M, K, N):
    rng = np.random.default_rng(M + K + N + t)
    W = refq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got, _ = harness.mul_mat(b200, t, W, X)
    ref, _ = harness.mul_mat(harness.cpu("x64"), t, W, X, n_threads=4)
    assert oracle.nmse(ref, got) <= 5e-4
    assert oracle.nmse(ref, got) <= (1e-10 if N <= 8 else 1e-6)   # more than 8 columns: fp16-operand prefill GEMM


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
def test_backend_mul_mat_id_matches_ggml_cpu(harness, refq, oracle, b200, t):
    rng = np.random.default_rng(t)
    n_as, M, K, n_used, n_tok = 8, 64, 512, 2, 7
    W = refq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    B = rng.uniform(-1, 1, (n_tok, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    got, _ = harness.mul_mat_id(b200, t, W, B, ids)
    ref, _ = harness.mul_mat_id(harness.cpu("x64"), t, W, B, ids, n_threads=4)
    assert oracle.nmse(ref, got) <= 1e-10


FUSED = "ggml_backend_b200_fused_epilogues"


@pytest.mark.parametrize("t,K", [(Q4_0, 512), (Q8_0, 1024), (Q4_K, 4096), (Q6_K, 2048)])
def test_backend_fused_ffn_graph(harness, refq, oracle, b200, t, K):
    """K7 through the reference's public API: {MUL_MAT [, ADD], MUL_MAT [, ADD], GLU} and {MUL_MAT, ADD} graphs. One column: the
    backend's peephole runs the whole pattern as one launch (its counter moves); three columns: the nodes run one by one."""
    rng = np.random.default_rng(K + t)
    M = 200
    Wu = refq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    Wg = refq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    bu, bg = rng.uniform(-1, 1, M).astype(np.float32), rng.uniform(-1, 1, M).astype(np.float32)
    cpu = harness.cpu("x64")
    for N in (1, 3):
        X = (rng.uniform(-1, 1, (N, K)) * 0.2).astype(np.float32)
        for op, bias in ((2, False), (2, True), (1, True), (3, False), (0, False), (4, True), (5, False), (-1, True)):
            args = (t, Wu, Wg if op >= 0 else None, X, bu if bias else None, bg if bias and op >= 0 else None)
            n0 = harness.proc_u64(b200, FUSED)
            got = harness.mul_mat_glu(b200, *args, glu_op=op, alpha=1.5, limit=4.0)
            n1 = harness.proc_u64(b200, FUSED)
            ref = harness.mul_mat_glu(cpu, *args, glu_op=op, alpha=1.5, limit=4.0, n_threads=4)
            assert oracle.nmse(ref, got) <= 1e-6, (op, bias, N)
            assert n1 - n0 == (1 if N == 1 else 0), (op, bias, N)


def test_backend_fused_moe_ffn_graph(harness, refq, oracle, b200):
    """the MUL_MAT_ID / ADD_ID forms (gpt-oss style experts with biases and SWIGLU_OAI): one token fused, several tokens node by node"""
    rng = np.random.default_rng(123)
    t, n_as, M, K, n_used = Q4_K, 8, 96, 1024, 4
    Wu = refq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    Wg = refq.quantize_weights(t, rng.uniform(-1, 1, (n_as * M, K)).astype(np.float32)).reshape(n_as, M, -1)
    bu, bg = rng.uniform(-1, 1, (n_as, M)).astype(np.float32), rng.uniform(-1, 1, (n_as, M)).astype(np.float32)
    cpu = harness.cpu("x64")
    for n_tok in (1, 5):
        B = (rng.uniform(-1, 1, (n_tok, 1, K)) * 0.2).astype(np.float32)
        ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
        for op, bias in ((3, True), (2, False), (-1, True)):
            args = (t, Wu, Wg if op >= 0 else None, B, ids, bu if bias else None, bg if bias and op >= 0 else None)
            n0 = harness.proc_u64(b200, FUSED)
            got = harness.mul_mat_id_glu(b200, *args, glu_op=op)
            n1 = harness.proc_u64(b200, FUSED)
            ref = harness.mul_mat_id_glu(cpu, *args, glu_op=op, n_threads=4)
            assert oracle.nmse(ref, got) <= 1e-6, (op, bias, n_tok)
            assert n1 - n0 == (1 if n_tok == 1 else 0), (op, bias, n_tok)


@pytest.mark.parametrize("op,min_tests", [("MUL_MAT_VEC_FUSION", 80), ("REGLU,GEGLU,SWIGLU,GEGLU_ERF,GEGLU_QUICK", 20), ("SWIGLU_OAI", 4), ("ADD", 20), ("MUL", 20),
                                          ("ADD_ID", 4)])
def test_reference_conformance_harness_k7(pkg, op, min_tests):
    """the reference's own cases for the K7 nodes (test_mul_mat_vec_fusion runs whole graphs: tests/test-backend-ops.cpp:5120-5249)"""
    if not have_ref() or not (REF_DIR / "test-backend-ops").exists():
        pytest.skip("oracle/_ref/test-backend-ops not built")
    rc, out = run_tbo(pkg, "test", "-b", "B2000", "-o", op)
    tail = "\n".join(out.splitlines()[-25:])
    m = re.search(r"(\d+)/(\d+) tests passed", out)
    assert m, tail
    assert m.group(1) == m.group(2), tail
    assert int(m.group(2)) >= min_tests, tail
    assert rc == 0, tail


@pytest.mark.parametrize("op,min_tests", [("RMS_NORM", 4), ("GET_ROWS", 20), ("ROPE", 40), ("SET_ROWS", 8), ("CPY", 20), ("CONT", 8), ("DUP", 4),
                                          ("FLASH_ATTN_EXT", 100)])
def test_reference_conformance_harness_f3(pkg, op, min_tests):
    """SURVEY 8(f) rank 3, first slice: the reference's own cases (tests/test-backend-ops.cpp) for the ops between the matmuls of a
    decode graph, each compared with ggml-cpu by the harness's own NMSE bar; what supports_op declines (other types, mrope,
    head sizes outside 64 / 128 / 256) is reported 'not supported' and does not count"""
    if not have_ref() or not (REF_DIR / "test-backend-ops").exists():
        pytest.skip("oracle/_ref/test-backend-ops not built")
    rc, out = run_tbo(pkg, "test", "-b", "B2000", "-o", op, timeout=1500)
    tail = "\n".join(out.splitlines()[-25:])
    m = re.search(r"(\d+)/(\d+) tests passed", out)
    assert m, tail
    failed = [ln for ln in out.splitlines() if "FAIL" in ln][:10]
    assert m.group(1) == m.group(2), "\n".join(failed) + "\n" + tail
    assert int(m.group(2)) >= min_tests, tail
    assert rc == 0, tail


def run_tbo(pkg, *args, timeout=900):
    env = dict(os.environ, GGML_BACKEND_PATH=str(pkg.BACKEND_PATH))
    p = subprocess.run([str(REF_DIR / "test-backend-ops"), *args], env=env, capture_output=True, text=True, timeout=timeout)
    return p.returncode, re.sub(r"\x1b\[[0-9;]*m", "", p.stdout + p.stderr)


@pytest.mark.parametrize("op", ["MUL_MAT", "MUL_MAT_ID"])
def test_reference_conformance_harness(pkg, op):
    """test-backend-ops test -b B2000 -o <op>: every case the backend claims through supports_op must pass the reference's
    own NMSE check against ggml-cpu; cases outside the hot path (f16/f32/other quants) are reported 'not supported'."""
    if not have_ref() or not (REF_DIR / "test-backend-ops").exists():
        pytest.skip("oracle/_ref/test-backend-ops not built")
    rc, out = run_tbo(pkg, "test", "-b", "B2000", "-o", op)
    tail = "\n".join(out.splitlines()[-25:])
    assert "Backend B2000: OK" in out or re.search(r"\d+/\d+ tests passed", out), tail
    m = re.search(r"(\d+)/(\d+) tests passed", out)
    assert m, tail
    assert m.group(1) == m.group(2), tail
    assert int(m.group(2)) > 50, tail   # it really ran our quantized cases
    assert rc == 0, tail


@pytest.mark.parametrize("t", [Q4_0, Q8_0, Q4_K])
@pytest.mark.parametrize("M,K,N,split", [(512, 1024, 1, None), (700, 512, 3, None), (1024, 2048, 40, None), (640, 1024, 2, [3.0, 1.0]),
                                         (100, 256, 1, None)])
def test_split_buffer_type_mul_mat(harness, refq, oracle, b200, t, M, K, N, split):
    """ggml_backend_split_buffer_type through the registry's get_proc_address (ggml-backend.h:199-200), the way llama places weights
    with --split-mode row (src/llama-model.cpp:385-403): the weight tensor's rows are spread over the devices of the process by the
    get_row_split rule, a MUL_MAT on it must match ggml-cpu, repeated computes must agree, and reading the tensor back must return
    the original bytes. On a one-GPU box the single device owns every row; with more devices the slices are real (M = 100 leaves
    devices without rows: boundaries are rounded down to 128)."""
    rng = np.random.default_rng(M + K + N + t)
    W = refq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    n_dev = harness.lib.refh_reg_dev_count(b200)
    ts = None if split is None else (split + [0.0] * 16)[:max(n_dev, 1)]
    if ts is not None and sum(ts) == 0:
        ts = None
    got, back = harness.mul_mat_split(b200, t, W, X, tensor_split=ts, repeats=3)
    ref, _ = harness.mul_mat(harness.cpu("x64"), t, W, X, n_threads=4)
    assert np.array_equal(back, W.reshape(back.shape))
    assert oracle.nmse(ref, got) <= (1e-10 if N <= 8 else 1e-6)
    plain, _ = harness.mul_mat(b200, t, W, X)
    if n_dev == 1:
        assert np.array_equal(plain, got)      # one device: the same kernels on the same rows
