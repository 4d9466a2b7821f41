#!/usr/bin/env python3
"""mmf_check.py — development check of the kind::f16 prefill GEMM (mmf16.cuh): NMSE against the oracle on the MMQ test shapes
and against the exact kind::i8 kernel at full size. The tile shape / cluster size come from GGML_B200_MMF_BN / _CL."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    import torch
    import b200_pkg
    from oracle.pyoracle import Oracle, RefQuant, Q4_0, Q8_0, Q4_K
    pkg = b200_pkg.load()
    oracle = Oracle()
    rq = RefQuant()
    shapes = [(128, 128, 128), (200, 256, 64), (130, 4096, 33), (512, 1024, 512), (64, 32, 16), (96, 96, 130), (257, 2048, 1), (300, 14336, 40),
              (1024, 512, 300), (640, 768, 257)]
    ctx = pkg.Context(0)
    worst = 0.0
    for t in (Q4_0, Q8_0, Q4_K):
        for (M, K, N) in shapes:
            if t == Q4_K and K % 256:
                continue
            rng = np.random.default_rng(M + 3 * K + 7 * N + t)
            W = rq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
            X = (rng.standard_normal((N, K)) * rng.choice([0.01, 1.0, 30.0, 1e6, 1e-6], (N, 1))).astype(np.float32)
            ctx.force_path(pkg.PATH_MMQ)
            ctx.mmq_f16(True)
            got = ctx.mul_mat(t, torch.from_numpy(W).cuda(), torch.from_numpy(X).cuda()).cpu().numpy()
            e = oracle.nmse(oracle.mul_mat(t, W, X), got)
            worst = max(worst, e)
            print(f"type {t} M={M} K={K} N={N}: nmse {e:.3e}", flush=True)
    # full size: f16 against i8 on the GPU
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    slabs = np.load(ROOT / "tests" / "golden" / "weight_slabs.npz")
    for name, t in (("q4_0", Q4_0), ("q8_0", Q8_0), ("q4_K", Q4_K)):
        slab = torch.from_numpy(slabs[name]).cuda()
        for (M, K, N) in ((4096, 4096, 512), (14336, 4096, 512), (4096, 14336, 512), (1024, 4096, 512)):
            idx = torch.randint(0, slab.shape[0], (M, K // 256), generator=g, device="cuda")
            W = slab[idx].reshape(M, -1).contiguous()
            X = (torch.randn((N, K), generator=g, device="cuda") * 3).contiguous()
            ctx.mmq_f16(False)
            a = ctx.mul_mat(t, W, X).float()
            ctx.mmq_f16(True)
            b = ctx.mul_mat(t, W, X).float()
            torch.cuda.synchronize()
            e = float(((a - b) ** 2).sum() / (a ** 2).sum())
            worst = max(worst, e)
            print(f"{name} M={M} K={K} N={N}: f16 vs i8 nmse {e:.3e}", flush=True)
    print(f"WORST {worst:.3e} BN={os.environ.get('GGML_B200_MMF_BN', 'auto')} CL={os.environ.get('GGML_B200_MMF_CL', 'auto')}")
    ctx.close()


if __name__ == "__main__":
    main()
