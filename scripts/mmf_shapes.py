#!/usr/bin/env python3
"""mmf_shapes.py — device time of one prefill MUL_MAT per Llama-3-8B shape (CUDA events on the launching stream, weight cache on):
where the prefill step's time goes, per shape and per arithmetic mode."""
import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_K")
    ap.add_argument("--tokens", default="512")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--i8", action="store_true")
    args = ap.parse_args()
    import torch
    import b200_pkg
    pkg = b200_pkg.load()
    dev = torch.device("cuda", 0)
    t = bench.TYPES[args.type]
    slab = torch.from_numpy(np.load(ROOT / "tests" / "golden" / "weight_slabs.npz")[args.type]).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    stream = torch.cuda.Stream(device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    with torch.cuda.stream(stream):
        ctx = pkg.Context(0)
        ctx.weight_cache(True)
        ctx.mmq_f16(not args.i8)
        tot = {}
        for (name, M, K, cnt) in (("wq", 4096, 4096, 2), ("wk", 1024, 4096, 2), ("gate", 14336, 4096, 2), ("down", 4096, 14336, 1), ("lm_head", 128256, 4096, 1 / 32)):
            idx = torch.randint(0, slab.shape[0], (M, K // 256), generator=g, device=dev)
            W = slab[idx].reshape(M, -1).contiguous()
            for N in [int(v) for v in args.tokens.split(",")]:
                X = torch.randn((N, K), generator=g, device=dev).contiguous()
                out = ctx.mul_mat(t, W, X)
                stream.synchronize()
                us = []
                for _ in range(args.reps):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    ctx.mul_mat(t, W, X, out=out)
                    e1.record(stream)
                    e1.synchronize()
                    us.append(e0.elapsed_time(e1) * 1e3)
                u = float(np.median(us))
                tot[N] = tot.get(N, 0.0) + u * cnt * 32
                print(f"{args.type} {name:8s} {M:6d}x{K:5d} N={N:5d}: {u:8.1f} us  {2.0 * M * K * N / u * 1e-6:7.1f} TFLOP/s", flush=True)
        for N, u in tot.items():
            print(f"model estimate N={N}: {u * 1e-3:.2f} ms  ({2.0 * N * 7.5e9 / u * 1e-6:.0f} TFLOP/s)")


if __name__ == "__main__":
    main()
