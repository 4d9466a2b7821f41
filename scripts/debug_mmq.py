import sys, numpy as np, torch
sys.path.insert(0, '.')
import b200_pkg
from oracle.pyoracle import Oracle, RefQuant, Q4_0, Q8_0, Q4_K
pkg = b200_pkg.load(); ctx = pkg.Context(0); orc = Oracle(); rq = RefQuant()
def run(t, M, K, N, reps=2):
    rng = np.random.default_rng(1)
    W = rq.quantize_weights(t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    ref = orc.mul_mat(t, W, X)
    Wd, Xd = torch.from_numpy(W).cuda(), torch.from_numpy(X).cuda()
    ctx.force_path(pkg.PATH_MMQ)
    outs = [ctx.mul_mat(t, Wd, Xd).cpu().numpy() for _ in range(reps)]
    ctx.force_path(pkg.PATH_AUTO)
    for i, got in enumerate(outs):
        err = np.abs(got - ref)
        bad = err > 1e-3 * np.abs(ref).max()
        print(f"t={t} M={M} K={K} N={N} rep{i}: nmse={orc.nmse(ref, got):.3e} bad={bad.mean():.3f} zeros={np.mean(got == 0):.3f} nan={np.isnan(got).mean():.3f}")
        if bad.any():
            # got is [N, M]; bad fraction per 32-row block of M and per column
            rb = bad.reshape(N, -1)[:, : (M // 32) * 32].reshape(N, -1, 32).mean(axis=(0, 2))
            cb = bad.mean(axis=1)
            print("   bad per 32-row block:", np.round(rb, 2)[:16])
            print("   bad per column     :", np.round(cb, 2)[:48])
            ratio = got[bad] / ref[bad]
            print("   got/ref on bad     :", np.round(ratio[:12], 3))
    print("   identical across reps:", all(np.array_equal(outs[0], o, equal_nan=True) for o in outs))
for (M, K, N) in [(256, 256, 48), (256, 512, 48), (256, 512, 128), (256, 1024, 48), (256, 1024, 128), (128, 640, 16), (256, 1024, 256), (128, 1024, 127)]:
    run(Q8_0, M, K, N)
run(Q4_K, 256, 1024, 48)
