"""GPU numerics tests of the sm_100a chain against the plain PyTorch fp32 oracle (ops/oracle.py).

Selection is compared bit-exactly given the kernel's own moments, moments to a relative tolerance
(SURVEY.md §2.10).  Multi-rank behaviour on one GPU is emulated by running each "rank's" send kernel to
completion before the receive kernel starts (never two kernels that wait on each other, see
B200_PROFILING.md)."""
import numpy as np
import pytest
import torch

from gaussiank_sgd_b200.ops import chain as ch
from gaussiank_sgd_b200.ops import native as N
from gaussiank_sgd_b200.ops import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _data(n, seed, kind="normal"):
    g = torch.Generator().manual_seed(seed)
    if kind == "normal":
        return torch.randn(n, generator=g) * 0.01
    if kind == "shifted":
        return torch.randn(n, generator=g) * 0.01 + 0.005
    if kind == "ties":
        return torch.randint(-3, 4, (n,), generator=g).float() * 0.25
    if kind == "laplace":
        u = (torch.rand(n, generator=g) - 0.5) * 0.9999
        return -0.01 * torch.sign(u) * torch.log1p(-2 * u.abs())
    if kind == "sparse":          # fewer non-zeros than k at density 0.001: the DGC sample threshold (0) is accepted
        x = torch.zeros(n)
        m = max(n // 2000, 1)
        x[torch.randperm(n, generator=g)[:m]] = torch.randn(m, generator=g)
        return x
    raise KeyError(kind)


def _run(comp, g, r, density, segments=None, cap=None, chunk=8192, use_residual=True):
    n = g.numel()
    G, R = g.to(DEV).clone(), r.to(DEV).clone()
    c = ch.CompressChain(G, R, segments or [[(0, n)]], compressor=comp, density=density, cap=cap, chunk=chunk,
                         use_residual=use_residual)
    sb = ch.SlotBuffer(1, c.cap, DEV)
    c.compress(ch.make_dst([sb.slot_ptr(0)], 0))
    torch.cuda.synchronize()
    idx, val, seq = sb.read_slot(0)
    return c, G.cpu(), R.cpu(), idx, val, seq


def test_library_loaded():
    lib = N.cuda_lib()
    assert b"sm_100a" in lib.gk_version()


@pytest.mark.parametrize("comp", ["gaussian", "gaussian2", "redsync", "redsynctrim"])
@pytest.mark.parametrize("kind", ["normal", "laplace", "shifted"])
@pytest.mark.parametrize("n,density", [(1000, 0.001), (269722, 0.001), (3000001, 0.004), (50001, 0.015625)])
def test_threshold_compressors_match_oracle(comp, kind, n, density):
    g, r = _data(n, 1, kind), _data(n, 2, kind) * 0.3
    c, G, R, idx, val, seq = _run(comp, g, r, density)
    acc = g + r
    st = c.read_stats()[0]
    m, s = oracle.moments(acc)
    assert abs(st["mean"] - m) <= 1e-5 * max(abs(m), s) and abs(st["std"] - s) <= 1e-5 * s
    if comp.startswith("gaussian"):
        # selection given the kernel's own moments must be bit-exact
        oi, thr = oracle.gaussian_select(acc, density, 3 if comp == "gaussian" else 5,
                                         mean=float(st["mean"]), std=float(st["std"]))
        assert np.float32(thr) == st["thr"]
    else:
        oi = oracle.select(comp, acc, density)
    oi = oi[:c.cap]
    assert seq == 1
    assert torch.equal(idx, oi)
    assert torch.equal(val, acc[oi])
    exp_res = acc.clone()
    exp_res[oi] = 0
    assert torch.equal(R, exp_res)
    assert torch.count_nonzero(G) == 0


@pytest.mark.parametrize("comp", ["topk", "mink"])
@pytest.mark.parametrize("kind", ["normal", "ties", "laplace"])
@pytest.mark.parametrize("n,density", [(1000, 0.001), (269722, 0.001), (3000001, 0.004), (50001, 0.25)])
def test_exact_topk_matches_oracle(comp, kind, n, density):
    g, r = _data(n, 3, kind), _data(n, 4, kind) * 0.3
    c, G, R, idx, val, seq = _run(comp, g, r, density)
    acc = g + r
    k = ch.compute_k(n, density)
    oi = oracle.topk_select(acc, k, smallest=(comp == "mink"))
    assert idx.numel() == k
    assert torch.equal(idx, oi)
    assert torch.equal(val, acc[oi])
    exp_res = acc.clone()
    exp_res[oi] = 0
    assert torch.equal(R, exp_res)
    # the histogram scratch must come back clean for the next call
    assert int(c.hist.abs().sum()) == 0


def test_repeated_calls_and_sequence_numbers():
    n = 500000
    c = None
    G = torch.zeros(n, device=DEV)
    R = torch.zeros(n, device=DEV)
    c = ch.CompressChain(G, R, [[(0, n)]], compressor="gaussian", density=0.001, cap=n)
    sb = ch.SlotBuffer(1, c.cap, DEV, parities=2)
    dst = ch.make_dst([sb.slot_ptr(0)], sb.parity_stride)
    r_ref = torch.zeros(n)
    for it in range(5):
        g = _data(n, 10 + it)
        G.copy_(g)
        c.compress(dst)
        torch.cuda.synchronize()
        idx, val, seq = sb.read_slot(0, parity=(it + 1) & 1)
        st = c.read_stats()[0]
        acc = g + r_ref
        oi, _ = oracle.gaussian_select(acc, 0.001, 3, mean=float(st["mean"]), std=float(st["std"]))
        assert seq == it + 1
        assert torch.equal(idx, oi) and torch.equal(val, acc[oi])
        r_ref = acc.clone()
        r_ref[oi] = 0
        assert torch.equal(R.cpu(), r_ref)
    ctrl = c.read_ctrl()
    assert ctrl["seq"] == 5 and ctrl["err"] == 0 and all(d == 0 for d in ctrl["done"])


def test_segmented_and_padded_layouts():
    sizes = [64, 1000, 4097, 300000, 7, 2048, 65537]
    offs, total = [], 0
    for s in sizes:
        offs.append(total)
        total += (s + 63) // 64 * 64
    g, r = torch.zeros(total), torch.zeros(total)
    for i, (o, s) in enumerate(zip(offs, sizes)):
        g[o:o + s] = _data(s, 20 + i)
        r[o:o + s] = _data(s, 40 + i) * 0.5
    # per-tensor mode (threshold=0): one launch, 7 segments
    segs = [[(o, s)] for o, s in zip(offs, sizes)]
    for comp in ["gaussian", "topk"]:
        c, G, R, idx, val, _ = _run(comp, g, r, 0.01, segments=segs, chunk=4096)
        stats = c.read_stats()
        exp = []
        for j, (o, s) in enumerate(zip(offs, sizes)):
            acc = g[o:o + s] + r[o:o + s]
            if comp == "gaussian":
                oi, _ = oracle.gaussian_select(acc, 0.01, 3, mean=float(stats[j]["mean"]), std=float(stats[j]["std"]))
            else:
                oi = oracle.topk_select(acc, ch.compute_k(s, 0.01))
            exp.append(oi + o)
        assert torch.equal(idx, torch.cat(exp)), comp
    # merged mode: one segment made of all tensors, padding excluded from n
    c, G, R, idx, val, _ = _run("topk", g, r, 0.01, segments=[[(o, s) for o, s in zip(offs, sizes)]], chunk=4096)
    flat = torch.cat([g[o:o + s] + r[o:o + s] for o, s in zip(offs, sizes)])
    assert int(c.seg_np[0]["n"]) == sum(sizes)
    oi = oracle.topk_select(flat, ch.compute_k(sum(sizes), 0.01))
    amap = torch.cat([torch.arange(o, o + s) for o, s in zip(offs, sizes)])
    assert torch.equal(idx, amap[oi])


def test_cap_overflow_stays_in_residual():
    n = 200000
    g = _data(n, 5, "shifted")
    c, G, R, idx, val, _ = _run("gaussian", g, torch.zeros(n), 0.001, cap=1024)
    st = c.read_stats()[0]
    oi, _ = oracle.gaussian_select(g, 0.001, 3, mean=float(st["mean"]), std=float(st["std"]))
    ctrl = c.read_ctrl()
    assert ctrl["total"] == oi.numel() and ctrl["sent"] == 1024 and ctrl["overflow"] == oi.numel() - 1024
    assert torch.equal(idx, oi[:1024])
    assert torch.equal(R[oi[1024:]], g[oi[1024:]])


def test_emulated_multi_rank_exchange_scatter_unscatter():
    """P 'ranks' send into one double-buffered arena (sequentially), then one receive kernel waits on
    the headers (already complete), scatter-adds the mean; unscatter restores the zero gradient."""
    n, P = 1000003, 8
    G = torch.zeros(n, device=DEV)
    R = torch.zeros(n, device=DEV)
    c = ch.CompressChain(G, R, [[(0, n)]], compressor="gaussian", density=0.001)
    arena = ch.SlotBuffer(P, c.cap, DEV, parities=2)
    for it in range(3):
        pairs = []
        seq0 = c.read_ctrl()["seq"]
        for r in range(P):
            G.copy_(_data(n, 100 * it + r))
            R.zero_()
            ctrl = c.read_ctrl()
            # every emulated rank is at the same sequence number
            c.ctrl.view(torch.int32)[10] = seq0
            c.compress(ch.make_dst([arena.slot_ptr(r)], arena.parity_stride))
            torch.cuda.synchronize()
            i, v, s = arena.read_slot(r, parity=(seq0 + 1) & 1)
            assert s == seq0 + 1
            pairs.append((i, v))
        dense = torch.zeros(n, device=DEV)
        ch.scatter(arena.src(wait=True, timeout_s=2.0), c.ctrl, dense, 1.0 / P)
        torch.cuda.synchronize()
        assert c.read_ctrl()["err"] == 0
        exp = oracle.decompress_mean(pairs, n, P)
        assert torch.equal(dense.cpu(), exp)     # rank-ordered adds: bit-identical to the sequential oracle
        ch.unscatter(arena.src(), c.ctrl, dense)
        torch.cuda.synchronize()
        assert torch.count_nonzero(dense) == 0


def test_receive_timeout_sets_error_flag_instead_of_hanging():
    n = 4096
    G = torch.zeros(n, device=DEV)
    c = ch.CompressChain(G, torch.zeros(n, device=DEV), [[(0, n)]], compressor="topk", density=0.01)
    arena = ch.SlotBuffer(2, c.cap, DEV, parities=2)
    G.copy_(_data(n, 1))
    c.compress(ch.make_dst([arena.slot_ptr(0)], arena.parity_stride))   # slot 1 never arrives
    dense = torch.zeros(n, device=DEV)
    ch.scatter(arena.src(wait=True, timeout_s=0.05), c.ctrl, dense, 0.5)
    torch.cuda.synchronize()
    assert c.read_ctrl()["err"] == 1
    assert torch.count_nonzero(dense) == 0


def test_randomk_gpu_matches_cpu_twin():
    n = 300000
    g = _data(n, 60)
    out = {}
    for dev in ("cpu", DEV):
        G, R = g.to(dev).clone(), torch.zeros(n, device=dev)
        c = ch.CompressChain(G, R, [[(0, n)]], compressor="topk", density=0.01)
        sb = ch.SlotBuffer(1, c.cap, dev)
        c.randomk_send(ch.make_dst([sb.slot_ptr(0)], 0), seed=77)
        if dev != "cpu":
            torch.cuda.synchronize()
        i, v, _ = sb.read_slot(0)
        out[dev] = (i, v, R.cpu())
    assert torch.equal(out["cpu"][0], out[DEV][0]) and torch.equal(out["cpu"][1], out[DEV][1])
    assert torch.equal(out["cpu"][2], out[DEV][2])
    assert out[DEV][0].unique().numel() == 3000


def test_fused_sgd_matches_torch_optim_bitwise():
    torch.manual_seed(0)
    shapes = [(64, 3, 7, 7), (64,), (1000, 512), (1000,), (3, 5)]
    ref = [torch.nn.Parameter(torch.randn(s, device=DEV)) for s in shapes]
    groups = [{"params": [ref[1], ref[3]], "weight_decay": 0.0}, {"params": [ref[0], ref[2], ref[4]], "weight_decay": 1e-4}]
    opt = torch.optim.SGD(groups, lr=0.05, momentum=0.875)
    offs, total = [], 0
    for p in ref:
        offs.append(total)
        total += (p.numel() + 63) // 64 * 64
    P, G, M = (torch.zeros(total, device=DEV) for _ in range(3))
    for o, p in zip(offs, ref):
        P[o:o + p.numel()] = p.detach().view(-1)
    tables = ch.FlatTables([(o, p.numel()) for o, p in zip(offs, ref)], [1, 0, 1, 0, 1], DEV, chunk=4096)
    hyp = [{"lr": 0.05, "momentum": 0.875, "weight_decay": 0.0}, {"lr": 0.05, "momentum": 0.875, "weight_decay": 1e-4}]
    for step in range(4):
        for o, q in zip(offs, ref):
            gr = torch.randn(q.shape, device=DEV)
            q.grad = gr.clone()
            G[o:o + q.numel()] = gr.view(-1)
        opt.step()
        ch.sgd_step(P, G, M, tables, ch.make_hyper(hyp, DEV, first_step=(step == 0)).to(DEV))
        for o, q in zip(offs, ref):
            assert torch.allclose(P[o:o + q.numel()], q.detach().view(-1), rtol=0, atol=1e-7), step


def test_lars_grad_norm_bucket_histogram():
    torch.manual_seed(1)
    sizes, offs, total = [300, 41000, 17], [0, 320, 41344], 41408
    P, G, A = torch.zeros(total, device=DEV), torch.zeros(total, device=DEV), torch.ones(total, device=DEV)
    ws, gs = [], []
    for o, s in zip(offs, sizes):
        w, g = torch.randn(s), torch.randn(s) * 0.1
        P[o:o + s], G[o:o + s] = w.to(DEV), g.to(DEV)
        ws.append(w); gs.append(g)
    tables = ch.FlatTables(list(zip(offs, sizes)), [0, 0, 0], DEV, chunk=4096)
    out = ch.grad_norm(G, tables, 0.25).cpu()
    tot = torch.cat(gs).norm(2)
    assert abs(out[0] - tot) < 1e-5 * tot and abs(out[1] - min(1.0, 0.25 / (tot + 1e-6))) < 1e-6
    hyp = [{"lr": 0.1, "momentum": 0.9, "weight_decay": 5e-4, "eeta": 0.001, "epsilon": 1e-5}]
    ch.lars_step(P, G, A, tables, ch.make_hyper(hyp, DEV, True).to(DEV))
    for o, s, w, g in zip(offs, sizes, ws, gs):
        w2, a2 = oracle.lars_reference(w, g, torch.ones(s), 0.1, 0.9, 5e-4, 0.001, 1e-5)
        assert torch.allclose(P[o:o + s].cpu(), w2, rtol=1e-5, atol=1e-7)
        assert torch.allclose(A[o:o + s].cpu(), a2, rtol=1e-5, atol=1e-7)
    # bucket compressor
    n = 50000
    g = _data(n, 50)
    Gb = g.to(DEV).clone()
    c = ch.CompressChain(Gb, torch.zeros(n, device=DEV), [[(0, n)]], compressor="topk", density=0.01)
    means = torch.zeros(2, device=DEV)
    c.bucket_means(means)
    assert torch.allclose(means.cpu(), torch.stack([g[g >= 0].mean(), g[g < 0].mean()]), rtol=1e-5)
    gm = torch.tensor([0.02, -0.03], device=DEV)
    c.bucket_reconstruct(means, gm)
    m = means.cpu()
    exp = torch.where(g >= 0, (g - m[0]) + 0.02, (g - m[1]) + (-0.03))
    assert torch.allclose(Gb.cpu(), exp, rtol=0, atol=1e-8)
    # histogram
    import ctypes as C
    x = torch.randn(100000, device=DEV)
    out_h = torch.zeros(64, dtype=torch.int32, device=DEV)
    N.check(N.cuda_lib().gk_histogram(x.data_ptr(), x.numel(), C.c_float(-4), C.c_float(4), 64, out_h.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream))
    ref_h = torch.histc(x.clamp(-3.9999, 3.9999), bins=64, min=-4, max=4)
    assert int(out_h.sum()) == 100000 and (out_h.cpu().float() - ref_h.cpu()).abs().max() <= 2


# --------------------------------------------------------------------------- round-2 kernels
def test_row_maxima_written_by_k1():
    n = 100003
    g, r = _data(n, 7, "laplace"), _data(n, 8) * 0.2
    G, R = g.to(DEV).clone(), r.to(DEV).clone()
    c = ch.CompressChain(G, R, [[(0, n)]], compressor="gaussian", density=0.001, chunk=4096)
    c.accumulate()
    torch.cuda.synchronize()
    acc = (g + r).abs()
    rows_per = c.rpb
    rm = c.rowmax.cpu().view(torch.int32)
    for e, (start, ln, _s, _o) in enumerate(c.blk_np):
        seg = acc[start:start + ln]
        nrow = (ln + 127) // 128
        pad = torch.zeros(nrow * 128)
        pad[:ln] = seg
        exp = pad.view(nrow, 128).max(dim=1).values.view(torch.int32)
        assert torch.equal(rm[e * rows_per: e * rows_per + nrow], exp), e


@pytest.mark.parametrize("kind", ["normal", "laplace", "sparse"])
@pytest.mark.parametrize("n,density", [(269722, 0.001), (3000001, 0.004), (50001, 0.009)])
def test_dgc_sampling_gpu_matches_cpu_twin(kind, n, density):
    g, r = _data(n, 11, kind), _data(n, 11, kind) * 0.3
    out = {}
    for dev in ("cpu", DEV):
        G, R = g.to(dev).clone(), r.to(dev).clone()
        c = ch.CompressChain(G, R, [[(0, n)]], compressor="dgcsampling", density=density)
        sb = ch.SlotBuffer(1, c.cap, dev)
        c.compress(ch.make_dst([sb.slot_ptr(0)], 0), seed=1234)
        if dev != "cpu":
            torch.cuda.synchronize()
            assert int(c.hist.abs().sum()) == 0
        i, v, _ = sb.read_slot(0)
        st = c.read_stats()[0]
        out[dev] = (i, v, R.cpu(), int(st["tbits"]), int(st["ties"]))
    assert out["cpu"][3:] == out[DEV][3:]                       # same threshold / tie budget
    assert torch.equal(out["cpu"][0], out[DEV][0]) and torch.equal(out["cpu"][1], out[DEV][1])
    assert torch.equal(out["cpu"][2], out[DEV][2])
    k = ch.compute_k(n, density)
    acc = g + r
    i = out[DEV][0]
    assert 0 < i.numel() <= 4 * k / 3 + 1
    if kind == "sparse":
        assert out[DEV][3] == 0 and i.numel() == int((acc != 0).sum())     # sampled threshold accepted
    # everything sent is at least as large as everything kept (threshold or exact top-k semantics)
    keep = torch.ones(n, dtype=torch.bool)
    keep[i] = False
    assert acc[i].abs().min() >= acc[keep].abs().max()


def _emulated_lists(n, P, k, seed, overlap=200):
    g = torch.Generator().manual_seed(seed)
    first = torch.randperm(n, generator=g)[:k].sort().values
    lists = []
    for r in range(P):
        sel = torch.randperm(n, generator=g)[:k]
        if r > 0:
            sel = torch.cat([sel, first[:overlap]])      # coordinates held by several ranks
        sel = sel.unique()                               # sorted ascending
        lists.append((sel, torch.randn(sel.numel(), generator=g)))
    return lists


def _fill_arena(arena, lists, seq):
    P = len(lists)
    for r, (sel, v) in enumerate(lists):
        words = sel.numpy().astype(np.uint64) | (v.numpy().view(np.uint32).astype(np.uint64) << np.uint64(32))
        base = ((seq & 1) * P + r) * arena.slot_stride
        arena.t[base] = int(sel.numel()) | (seq << 32)
        arena.t[base + N.GK_SLOT_HDR: base + N.GK_SLOT_HDR + sel.numel()] = torch.from_numpy(words.view(np.int64)).to(DEV)


@pytest.mark.parametrize("P", [1, 2, 8])
@pytest.mark.parametrize("momentum,nesterov", [(0.9, False), (0.9, True), (0.0, False)])
def test_folded_receive_update_matches_materialised_path(P, momentum, nesterov):
    """gk_sgd_recv (no dense gradient) against gk_scatter + gk_sgd on the same pair lists: identical on the
    untouched coordinates, within the rounding of the split FMA on the received ones; the arena used as the
    rank bitmask comes back all zero; two runs give the same bits (rank-ordered sums)."""
    torch.manual_seed(5)
    sizes = [300, 410000, 17, 50000, 1000003]
    offs, total = [], 0
    for s in sizes:
        offs.append(total)
        total += (s + 63) // 64 * 64
    tables = ch.FlatTables(list(zip(offs, sizes)), [0, 1, 0, 1, 1], DEV)
    hyp = [{"lr": 0.05, "momentum": momentum, "weight_decay": 0.0, "nesterov": nesterov},
           {"lr": 0.1, "momentum": momentum, "weight_decay": 1e-3, "nesterov": nesterov}]
    valid = torch.cat([torch.arange(o, o + s) for o, s in zip(offs, sizes)])
    P0, M0 = torch.randn(total, device=DEV), torch.randn(total, device=DEV) * 0.1
    ctrl = torch.zeros(ch.CTRL_DTYPE.itemsize, dtype=torch.uint8, device=DEV)
    for step, first in enumerate((True, False)):
        lists = [(valid[s], v) for s, v in _emulated_lists(valid.numel(), P, 3000, 40 + step)]
        cap = max(l[0].numel() for l in lists) + 8
        arena = ch.SlotBuffer(P, cap, DEV, parities=2)
        seq = 7 + step
        ctrl.view(torch.int32)[10] = seq
        _fill_arena(arena, lists, seq)
        hy = ch.make_hyper(hyp, DEV, first).to(DEV)
        mom = lambda M: M if momentum else None   # noqa: E731
        P1, M1, G = P0.clone(), M0.clone(), torch.zeros(total, device=DEV)
        ch.scatter(arena.src(wait=True, timeout_s=2.0), ctrl, G, 1.0 / P)
        ch.sgd_step(P1, G, mom(M1), tables, hy)
        res = []
        for _ in range(2):
            P2, M2, mask = P0.clone(), M0.clone(), torch.zeros(total, device=DEV)
            ch.sgd_recv_step(P2, mask.view(torch.int32), mom(M2), tables, hy, arena.src(wait=True, timeout_s=2.0),
                             ctrl, 1.0 / P)
            torch.cuda.synchronize()
            assert torch.count_nonzero(mask) == 0
            res.append((P2, M2))
        assert int(ctrl.view(torch.int32)[11]) == 0
        assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
        P2, M2 = res[0]
        touched = torch.zeros(total, dtype=torch.bool, device=DEV)
        for sel, _ in lists:
            touched[sel.to(DEV)] = True
        assert torch.equal(P1[~touched], P2[~touched]) and torch.equal(M1[~touched], M2[~touched])
        assert (P1 - P2).abs().max() <= 4 * torch.finfo(torch.float32).eps * max(1.0, float(P1.abs().max()))
        assert (M1 - M2).abs().max() <= 4 * torch.finfo(torch.float32).eps * max(1.0, float(M1.abs().max()))
        P0, M0 = P1, M1


def test_folded_update_flags_unsorted_lists():
    n = 4096 * 4
    tables = ch.FlatTables([(0, n)], [0], DEV)
    hy = ch.make_hyper([{"lr": 0.1, "momentum": 0.9, "weight_decay": 0.0}], DEV, False).to(DEV)
    ctrl = torch.zeros(ch.CTRL_DTYPE.itemsize, dtype=torch.uint8, device=DEV)
    ctrl.view(torch.int32)[10] = 2
    arena = ch.SlotBuffer(2, 64, DEV, parities=2)
    a = (torch.tensor([5, 9, 100]), torch.ones(3))
    b = (torch.tensor([100, 9, 5]), torch.ones(3))           # descending: the owner's binary search must fail loudly
    _fill_arena(arena, [a, b], 2)
    Pm, Mm, mask = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    ch.sgd_recv_step(Pm, mask.view(torch.int32), Mm, tables, hy, arena.src(wait=True, timeout_s=1.0), ctrl, 0.5)
    torch.cuda.synchronize()
    assert int(ctrl.view(torch.int32)[11]) == 2


def test_gaussian_heavy_tail_needs_several_rounds_and_matches_oracle():
    """A few huge entries inflate sigma, the first threshold selects far fewer than 2k/3 and the search
    descends (compression.py:375-376): the lazy rounds must follow it exactly."""
    n = 2000000
    g = _data(n, 21) * 0.1
    g[torch.randperm(n, generator=torch.Generator().manual_seed(3))[:40]] = 50.0
    c, G, R, idx, val, seq = _run("gaussian", g, torch.zeros(n), 0.001)
    st = c.read_stats()[0]
    oi, thr = oracle.gaussian_select(g, 0.001, 3, mean=float(st["mean"]), std=float(st["std"]))
    assert int(st["chosen"]) != 0                    # not settled by the first candidate
    assert np.float32(thr) == st["thr"] and torch.equal(idx, oi) and torch.equal(val, g[oi])
