"""b200_mul_mat_noalloc (include/ggml_b200.h): the decode MUL_MAT with its activation scratch in a static __device__ array - no
cudaMalloc / cudaFree on the call. No counterpart in the reference, so the check is
against the existing entry point b200_mul_mat on the same inputs, through the C-ABI, on two calls in a row (the scratch keeps the
previous call's records). Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest


def test_noalloc_symbol_and_limit(pkg):
    L = pkg.lib()
    assert hasattr(L, "b200_mul_mat_noalloc")
    assert L.b200_mul_mat_noalloc_max_units() == 1024


@pytest.mark.gpu
def test_noalloc_matches_mul_mat_two_calls_in_a_row(pkg, oracle, refq):
    import torch
    from oracle.pyoracle import Q4_0, Q4_K, Q8_0
    L = pkg.lib()
    ctx = pkg.Context(0)
    try:
        rng = np.random.default_rng(2024)
        # two calls in a row with different shapes, types and column counts: the second must not see the first's records
        cases = [(Q4_K, 640, 4096, 1), (Q4_0, 384, 2048, 2), (Q8_0, 256, 1024, 5), (Q4_K, 512, 8192, 8), (Q4_K, 640, 4096, 1)]
        for (t, M, K, N) in cases:
            W = torch.from_numpy(refq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))).cuda()
            Xh = rng.uniform(-1, 1, (N, K)).astype(np.float32)
            X = torch.from_numpy(Xh).cuda()
            want = ctx.mul_mat(t, W, X)
            torch.cuda.synchronize()
            d = pkg.MulMatDesc.contiguous(t, M, K, N)
            for rep in range(2):
                got = torch.full((N, M), float("nan"), dtype=torch.float32, device="cuda")
                free0 = torch.cuda.mem_get_info()[0]
                pkg.check(L.b200_mul_mat_noalloc(ctx.h, C.byref(d), W.data_ptr(), X.data_ptr(), got.data_ptr()), "b200_mul_mat_noalloc")
                ctx.sync()
                if rep == 1:   # (the first call of a shape may make the driver load a kernel image lazily; the second cannot)
                    assert torch.cuda.mem_get_info()[0] >= free0, "device memory was allocated during the call"
                err = oracle.nmse(want.cpu().numpy(), got.cpu().numpy())
                assert err <= 1e-10, (t, M, K, N, rep, err)      # the bound the decode parity tests use (tests/test_gpu_parity.py)
                ref = oracle.mul_mat(t, W.cpu().numpy(), Xh)
                assert oracle.nmse(ref, got.cpu().numpy()) <= 1e-10
        # beyond the static scratch: rejected with a clear error, nothing launched
        big = pkg.MulMatDesc.contiguous(Q4_K, 16, 256 * 200, 8)   # 8 x 200 = 1600 units > 1024
        l0 = ctx.launches
        st = L.b200_mul_mat_noalloc(ctx.h, C.byref(big), 256, 256, 256)
        assert st == 1 and b"static scratch" in L.b200_last_error() and ctx.launches == l0
        nine = pkg.MulMatDesc.contiguous(Q4_K, 16, 256, 9)
        assert L.b200_mul_mat_noalloc(ctx.h, C.byref(nine), 256, 256, 256) == 1
    finally:
        ctx.close()
