"""Row-split tensor parallelism: the host-side partition rule on CPU, the slice/gather layout with two gloo ranks,
and - on a GPU - the fused exchange epilogue of the kernels with a second destination on the same device."""
import os
import socket

import numpy as np
import pytest

from oracle.pyoracle import Q4_0, Q4_K, Q8_0


# ---------------------------------------------------------------- the partition rule (reference: get_row_split, ggml-cuda.cu:786-799)

def ref_row_split(nrows, tensor_split, dev, n_dev, rounding):
    """literal restatement of the reference function for the checker"""
    lo = 0 if dev == 0 else int(nrows * tensor_split[dev])
    lo -= lo % rounding
    if dev == n_dev - 1:
        hi = nrows
    else:
        hi = int(nrows * tensor_split[dev + 1])
        hi -= hi % rounding
    return lo, hi


@pytest.mark.parametrize("nrows", [1, 64, 127, 128, 1000, 4096, 14336, 128256])
@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_bounds_cover_disjoint_ordered(pkg, nrows, n):
    prev = 0
    for r in range(n):
        lo, hi = pkg.row_split_bounds(nrows, n, r)
        assert lo == prev and hi >= lo
        if r < n - 1:
            assert hi % pkg.ROW_SPLIT_ROUNDING == 0
        prev = hi
    assert prev == nrows


def test_bounds_follow_the_reference_rule(pkg):
    rng = np.random.default_rng(5)
    for _ in range(200):
        n = int(rng.integers(1, 9))
        w = rng.uniform(0.0, 1.0, n)
        if w.sum() == 0:
            continue
        cum = np.concatenate([[0.0], np.cumsum(w / w.sum())[:-1]]).tolist()   # llama.cpp's normalised cumulative form
        nrows = int(rng.integers(1, 200000))
        rounding = int(rng.choice([1, 32, 64, 128]))
        for r in range(n):
            assert pkg.row_split_bounds(nrows, n, r, cum, rounding) == ref_row_split(nrows, cum, r, n, rounding)


def test_bounds_reject_bad_splits(pkg):
    with pytest.raises(ValueError):
        pkg.row_split_bounds(100, 2, 0, [0.1, 0.5])
    with pytest.raises(ValueError):
        pkg.row_split_bounds(100, 3, 0, [0.0, 0.7, 0.5])


# ---------------------------------------------------------------- two ranks over gloo: slices + all-gather == the whole product

def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, t, M, K, N, split, q):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    import torch
    import torch.distributed as dist
    import b200_pkg
    from oracle.pyoracle import Oracle
    pkg = b200_pkg.load()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle = Oracle()
        rng = np.random.default_rng(11)                      # same seed on every rank: replicated weights / activations
        Wf = rng.uniform(-1, 1, (M, K)).astype(np.float32)
        X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
        W = oracle.quantize("q4_0" if t == Q4_0 else "q8_0", Wf)
        lo, hi = pkg.row_split_bounds(M, world, rank, split, 32)
        mine = oracle.mul_mat(t, np.ascontiguousarray(W[lo:hi]), X) if hi > lo else np.zeros((N, 0), np.float32)   # this rank's dst rows
        # the layout the fused all-gather produces: every rank's slice at columns [lo, hi) of a full [N, M] dst
        sizes = [pkg.row_split_bounds(M, world, r, split, 32) for r in range(world)]
        parts = [torch.zeros((N, h - l), dtype=torch.float32) for (l, h) in sizes]
        # gloo's all_gather wants equal shapes: pad to the widest slice
        wmax = max(h - l for (l, h) in sizes)
        padded = torch.zeros((N, wmax), dtype=torch.float32)
        padded[:, :hi - lo] = torch.from_numpy(np.ascontiguousarray(mine))
        gathered = [torch.zeros((N, wmax), dtype=torch.float32) for _ in range(world)]
        dist.all_gather(gathered, padded)
        for r, (l, h) in enumerate(sizes):
            parts[r] = gathered[r][:, :h - l]
        full = torch.cat(parts, dim=1).numpy()
        whole = oracle.mul_mat(t, W, X)
        q.put((rank, bool(np.array_equal(full, whole)), (lo, hi)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("split", [None, [0.0, 0.75], [0.0, 0.0]])
def test_two_rank_slices_gather_to_the_whole_product(split):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    M, K, N = 200, 256, 3
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, Q8_0, M, K, N, split, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for (_, ok, _) in res), res
    bounds = dict((r, b) for (r, _, b) in res)
    assert bounds[0][1] == bounds[1][0] and bounds[0][0] == 0 and bounds[1][1] == M


# ---------------------------------------------------------------- GPU: the fused exchange epilogue

@pytest.mark.gpu
@pytest.mark.parametrize("deferred", [False, True])
@pytest.mark.parametrize("t,N", [(Q4_K, 1), (Q4_0, 2), (Q8_0, 1), (Q4_K, 5), (Q4_K, 40), (Q8_0, 64)])
def test_fused_exchange_writes_every_destination(pkg, oracle, t, N, deferred):
    """rank 0 of a 2-rank split whose 'peer' dst is a second buffer on the same GPU and whose peer flag is already at the far
    epoch (nothing spins): the kernel must store every value into both destinations, bump its epoch, reset its CTA counter and
    publish the epoch into the peer's flag slot. Mat-vec (N <= 8) and tensor-core (N > 8) kernels."""
    import torch
    from oracle.pyoracle import RefQuant, have_ref

    def cu(a):
        return torch.from_numpy(np.ascontiguousarray(a)).cuda()

    def quantize_weights(oracle, t, w):
        if have_ref():
            return RefQuant().quantize_weights(t, w)
        if t == Q4_K:
            pytest.skip("q4_K weights need oracle/_ref")
        return oracle.quantize("q4_0" if t == Q4_0 else "q8_0", w)

    NMSE_TIGHT = 1e-10
    rng = np.random.default_rng(23 + t + N)
    M, K = 384, 1024
    W = quantize_weights(oracle, t, rng.uniform(-1, 1, (M, K)).astype(np.float32))
    X = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    c = pkg.Context(0)
    try:
        ref = c.mul_mat(t, cu(W), cu(X)).cpu().numpy()
        pool = torch.zeros(2 * N * M + 64, dtype=torch.float32, device="cuda")      # [dst of rank 0][dst of the 'peer'][sync words x2]
        sync = pool[2 * N * M:].view(torch.int32)
        sync[1] = 0x7FFFFFF0                                                     # rank 1's flag in rank 0's words: already there
        torch.cuda.synchronize()
        base = pool.data_ptr()
        peer = base + 4 * N * M
        s0, s1 = base + 8 * N * M, base + 8 * N * M + 128
        c.row_split(0, [base, peer], [s0, s1])
        c.row_split_deferred(deferred)       # deferred: the wait for the peers is a separate step (next kernel / row_split_wait)
        out = pool[:N * M].view(N, M)
        for rep in range(3):
            c.mul_mat(t, cu(W), cu(X), out=out)
            c.row_split_wait()
            torch.cuda.synchronize()
            assert np.array_equal(out.cpu().numpy(), ref)
            assert np.array_equal(pool[N * M:2 * N * M].view(N, M).cpu().numpy(), ref)
            words = pool[2 * N * M:].view(torch.int32).cpu().numpy()
            assert words[8] == rep + 1 and words[9] == 0          # epoch, CTA counter
            assert words[32 + 0] == rep + 1                        # published into the peer's slot for rank 0
        c.row_split(0, [base], [s0])                               # one rank: off
        assert np.array_equal(c.mul_mat(t, cu(W), cu(X)).cpu().numpy(), ref)
    finally:
        c.close()
    assert oracle.nmse(oracle.mul_mat(t, W, X), ref) <= (NMSE_TIGHT if N <= 8 else 1e-6)   # N > 8: fp16-operand prefill GEMM


@pytest.mark.gpu
@pytest.mark.parametrize("N", [1, 40])
def test_empty_row_slice_still_takes_part_in_the_exchange(pkg, N):
    """get_row_split can leave a rank without rows (M = 512 over 8 ranks, rounding 128). Such a rank must still count the launch:
    flag protocol - epoch bumped, CTA counter reset, flag published to the peers (the peers would spin and trap)."""
    import torch
    c = pkg.Context(0)
    try:
        pool = torch.zeros(2 * N * 64 + 64, dtype=torch.float32, device="cuda")
        sync = pool[2 * N * 64:].view(torch.int32)
        sync[1] = 0x7FFFFFF0
        torch.cuda.synchronize()
        base = pool.data_ptr()
        s0, s1 = base + 8 * N * 64, base + 8 * N * 64 + 128
        c.row_split(0, [base, base + 4 * N * 64], [s0, s1])
        d = pkg.MulMatDesc.contiguous(Q4_K, 0, 1024, N)          # this rank's slice: zero rows of a 64-row matrix
        d.nb1, d.nb2, d.nb3 = 4 * 64, 4 * 64 * N, 4 * 64 * N
        x = torch.zeros((N, 1024), dtype=torch.float32, device="cuda")
        for rep in range(3):
            l0 = c.launches
            pkg.check(pkg.lib().b200_mul_mat(c.h, __import__("ctypes").byref(d), 256, x.data_ptr(), base), "b200_mul_mat")
            c.sync()
            words = pool[2 * N * 64:].view(torch.int32).cpu().numpy()
            assert c.launches == l0 + 1
            assert words[8] == rep + 1 and words[9] == 0 and words[32] == rep + 1
        assert not pool[:2 * N * 64].any()
    finally:
        c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("t,N,graph", [(Q4_K, 1, False), (Q4_K, 1, True), (Q4_0, 2, False), (Q8_0, 1, False), (Q4_K, 5, True), (Q4_0, 8, False)])
def test_flagged_payload_exchange_loopback(pkg, oracle, t, N, graph):
    """rank 0 of a 2-rank split on ONE GPU with the flagged-payload exchange (b200_ctx_set_row_split_ll). The 'peer' staging area is
    this rank's own, shifted by the byte distance between a row of rank 0 and the same row of rank 1 - so the {value, epoch} records
    rank 0 sends land exactly where it polls for rank 1's rows: after every launch the lower half of the full dst must equal the
    upper half, bit for bit, over several launches (both epoch parities) and through CUDA-graph replays (epochs live on the device)."""
    import torch
    from oracle.pyoracle import RefQuant, have_ref

    def cu(a):
        return torch.from_numpy(np.ascontiguousarray(a)).cuda()

    if not have_ref() and t == Q4_K:
        pytest.skip("q4_K weights need oracle/_ref")
    rng = np.random.default_rng(77 + t + N)
    Ml, K = 384, 1024
    Mf = 2 * Ml
    w = rng.uniform(-1, 1, (Ml, K)).astype(np.float32)
    W = RefQuant().quantize_weights(t, w) if have_ref() else oracle.quantize("q4_0" if t == Q4_0 else "q8_0", w)
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        c = pkg.Context(0)                      # bound to `stream` (graph capture needs a non-default stream)
        try:
            Wd = cu(W)
            region = N * Mf                                                         # floats: the full dst [N][Mf]
            pool = torch.zeros(region + 64, dtype=torch.float32, device="cuda")     # [full dst][sync words x2]
            stage = torch.zeros(4 * region + 2 * Ml, dtype=torch.float32, device="cuda")
            xs = [cu(rng.uniform(-1, 1, (N, K)).astype(np.float32)) for _ in range(5)]
            refs = [c.mul_mat(t, Wd, xv).cpu().numpy() for xv in xs]                # the same kernel without the split
            base = pool.data_ptr()
            s0, s1 = base + 4 * region, base + 4 * region + 128
            c.row_split(0, [base, base], [s0, s1])
            c.row_split_ll([stage.data_ptr(), stage.data_ptr() + 8 * Ml], base, 4 * region, rounding=1)
            out = pool[:region].view(N, Mf)[:, :Ml]
            d = pkg.MulMatDesc.contiguous(t, Ml, K, N)
            d.nb1, d.nb2, d.nb3 = 4 * Mf, 4 * Mf * N, 4 * Mf * N
            x = torch.empty_like(xs[0])
            run = lambda: c.mul_mat(t, Wd, x, out=out, desc=d)   # noqa: E731
            if graph:
                x.copy_(xs[0])
                run()
                stream.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    run()
                run = g.replay
            for rep, xv in enumerate(xs):
                x.copy_(xv)
                pool[:region].zero_()
                run()
                stream.synchronize()
                full = pool[:region].view(N, Mf).cpu().numpy()
                assert np.array_equal(full[:, :Ml], refs[rep]), f"own rows differ at launch {rep}"
                assert np.array_equal(full[:, Ml:], refs[rep]), f"received rows differ at launch {rep}"
            words = pool[region:].view(torch.int32).cpu().numpy()
            assert words[9] == 0 and words[8] >= len(xs)                            # CTA counter reset, epoch counted on the device
            assert words[32] == 0                                                   # no flag was published: the data carried the epoch
        finally:
            c.close()


# ---------------------------------------------------------------- two real devices (skipped on a one-GPU box)

@pytest.mark.gpu
def test_flagged_payload_exchange_on_two_devices_full_70b_step(pkg, oracle):
    """Both ranks of a 2-way row split in ONE process on two real GPUs (peer access over NVLink): every MUL_MAT of a Llama-3-70B decode
    step (80 layers x 7 + lm_head = 561 ops, 401 launches per rank after the shared-src1 fusion) through the flagged-payload exchange.
    After each op BOTH ranks must hold the full dst, bit-identical to the same MUL_MAT computed unsplit on one GPU: exercises the epoch
    counting, both staging halves and who-owns-which-rows over several hundred dependent launches. Weights: one matrix per shape
    (shared by the layers), activations differ per op."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import bench
    L = pkg.lib()
    pkg.check(L.b200_enable_peer_access(0, 1), "peer 0->1")
    pkg.check(L.b200_enable_peer_access(1, 0), "peer 1->0")
    t, tname = Q4_K, "q4_K"
    m = bench.MODELS["70b"]
    shapes = dict((name, (M, K, grp)) for (name, M, K, grp) in bench.layer_ops(m))
    shapes["lm_head"] = (m["n_vocab"], m["n_embd"], "head_in")
    slab = bench.load_slab(tname)
    rng = np.random.default_rng(70)
    dev = [torch.device("cuda", 0), torch.device("cuda", 1)]
    Wfull, Wpart, bounds = {}, [{}, {}], {}
    for name, (M, K, _) in shapes.items():
        w = torch.from_numpy(slab[rng.integers(0, slab.shape[0], (M, K // 256))].reshape(M, -1))
        Wfull[name] = w.to(dev[0])
        bounds[name] = [pkg.row_split_bounds(M, 2, r, rounding=1) for r in range(2)]
        for r in range(2):
            lo, hi = bounds[name][r]
            Wpart[r][name] = w[lo:hi].contiguous().to(dev[r])
    order = []
    for _ in range(m["n_layer"]):
        order += [["wq", "wk", "wv"], ["wo"], ["gate", "up"], ["down"]]
    order.append(["lm_head"])
    n_out = sum(shapes[n][0] for grp in order for n in grp)
    pools, stages, ctxs, streams = [], [], [], []
    for r in range(2):
        with torch.cuda.device(r):
            pools.append(torch.zeros(n_out + 64, dtype=torch.float32, device=dev[r]))
            stages.append(torch.zeros(4 * n_out, dtype=torch.float32, device=dev[r]))
            streams.append(torch.cuda.Stream(device=dev[r]))
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    for r in range(2):
        with torch.cuda.device(r), torch.cuda.stream(streams[r]):
            c = pkg.Context(r)
            c.row_split(r, [pools[0].data_ptr(), pools[1].data_ptr()], [pools[0].data_ptr() + 4 * n_out, pools[1].data_ptr() + 4 * n_out])
            c.row_split_ll([stages[0].data_ptr(), stages[1].data_ptr()], pools[r].data_ptr(), 4 * n_out, rounding=1)
            ctxs.append(c)
    ref_ctx = None
    try:
        gx = torch.Generator(device=dev[0]); gx.manual_seed(5)
        xs = []
        for grp in order:   # all activations up front, on both devices (nothing but launches inside the exchange loop)
            x0 = (torch.rand((1, shapes[grp[0]][1]), generator=gx, device=dev[0]) * 2 - 1).contiguous()
            xs.append([x0, x0.to(dev[1])])
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        off, placed = 0, []
        for gi, grp in enumerate(order):
            K = shapes[grp[0]][1]
            offs = []
            for n in grp:
                offs.append(off)
                off += shapes[n][0]
            placed.append(offs)
            for r in range(2):
                with torch.cuda.device(r), torch.cuda.stream(streams[r]):
                    outs = [pools[r][o + bounds[n][r][0]:o + bounds[n][r][1]].view(1, -1) for n, o in zip(grp, offs)]
                    if len(grp) == 1:
                        n = grp[0]
                        d = pkg.MulMatDesc.contiguous(t, bounds[n][r][1] - bounds[n][r][0], K, 1)
                        d.nb1 = d.nb2 = d.nb3 = 4 * shapes[n][0]
                        ctxs[r].mul_mat(t, Wpart[r][n], xs[gi][r], out=outs[0], desc=d)
                    else:
                        ctxs[r].mul_mat_shared(t, [Wpart[r][n] for n in grp], xs[gi][r], outs=outs, dst_strides=[4 * shapes[n][0] for n in grp])
        for r in range(2):
            with torch.cuda.device(r):
                streams[r].synchronize()
        # the unsplit reference for a sample of the ops (every shape; first, middle and last layers), computed afterwards on one GPU
        n_checked = 0
        with torch.cuda.device(0):
            ref_ctx = pkg.Context(0)
            for gi, grp in enumerate(order):
                if not (gi % 40 == 0 or gi % 40 == 1 or gi % 40 == 2 or gi % 40 == 3 or gi >= len(order) - 2):
                    continue
                for n, o in zip(grp, placed[gi]):
                    ref = ref_ctx.mul_mat(t, Wfull[n], xs[gi][0]).view(-1).cpu()
                    for r in range(2):
                        got = pools[r][o:o + shapes[n][0]].cpu()
                        assert torch.equal(got, ref), f"op {gi} ({n}) on rank {r} differs from the unsplit result"
                    n_checked += 1
        assert n_checked >= 50
        for r in range(2):
            words = pools[r][n_out:].view(torch.int32).cpu().numpy()
            assert words[8] == len(order) and words[9] == 0     # one epoch per launch, CTA counter back at zero
    finally:
        if ref_ctx is not None:
            ref_ctx.close()
        for c in ctxs:
            c.close()
