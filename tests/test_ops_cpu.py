"""CPU tests of the selection rules: PyTorch oracle vs the reference's own compression.py (when the
reference tree is mounted) and C++ CPU twin vs the oracle (SURVEY.md §4 plan items 1-2)."""
import numpy as np
import pytest
import torch

from gaussiank_sgd_b200.ops import chain as ch
from gaussiank_sgd_b200.ops import oracle

COMPS = ["gaussian", "gaussian2", "topk", "mink", "redsync", "redsynctrim"]


def _data(n, seed, kind="normal"):
    g = torch.Generator().manual_seed(seed)
    if kind == "normal":
        return torch.randn(n, generator=g) * 0.01
    if kind == "laplace":
        u = (torch.rand(n, generator=g) - 0.5) * 0.9999
        return -0.01 * torch.sign(u) * torch.log1p(-2 * u.abs())
    if kind == "shifted":
        return torch.randn(n, generator=g) * 0.01 + 0.005
    if kind == "ties":
        return torch.randint(-3, 4, (n,), generator=g).float() * 0.25
    raise KeyError(kind)


def _run_cpu_chain(comp, g, r, density, segments=None, cap=None, chunk=8192, use_residual=True):
    n = g.numel()
    G, R = g.clone(), r.clone()
    c = ch.CompressChain(G, R, segments or [[(0, n)]], compressor=comp, density=density, cap=cap,
                         chunk=chunk, use_residual=use_residual)
    sb = ch.SlotBuffer(1, c.cap, "cpu")
    c.compress(ch.make_dst([sb.slot_ptr(0)], 0))
    idx, val, seq = sb.read_slot(0)
    return c, G, R, idx, val, seq


@pytest.mark.parametrize("comp", COMPS)
@pytest.mark.parametrize("kind", ["normal", "laplace", "shifted", "ties"])
@pytest.mark.parametrize("n,density", [(1000, 0.001), (269722, 0.001), (50001, 0.015625)])
def test_cpu_twin_matches_oracle(comp, kind, n, density):
    g, r = _data(n, 1, kind), _data(n, 2, kind) * 0.3
    c, G, R, idx, val, seq = _run_cpu_chain(comp, g, r, density)
    acc, oi, ov, nr = oracle.compress(comp, g, r, density, cap=c.cap)
    assert seq == 1
    assert torch.equal(idx, oi)
    assert torch.equal(val, ov)
    assert torch.equal(R, nr)
    assert torch.count_nonzero(G) == 0  # gradient consumed and zeroed


@pytest.mark.parametrize("comp", ["gaussian", "topk"])
def test_conservation_and_order(comp):
    """selected + residual == grad + old residual exactly; indices unique and ascending."""
    n = 77777
    g, r = _data(n, 3), _data(n, 4)
    c, G, R, idx, val, _ = _run_cpu_chain(comp, g, r, 0.004)
    acc = g + r
    rebuilt = R.clone()
    rebuilt[idx] += val
    assert torch.equal(rebuilt, acc)
    assert torch.all(idx[1:] > idx[:-1])
    if comp == "topk":
        assert idx.numel() == ch.compute_k(n, 0.004)
        assert acc.abs()[idx].min() >= R.abs().max()


def test_cap_keeps_overflow_in_residual():
    n = 20000
    g, r = _data(n, 5, "shifted"), torch.zeros(n)
    c, G, R, idx, val, _ = _run_cpu_chain("gaussian", g, r, 0.001, cap=1024)
    ctrl = c.read_ctrl()
    oi = oracle.select("gaussian", g, 0.001)
    assert ctrl["total"] == oi.numel()
    assert ctrl["sent"] == min(1024, oi.numel())
    assert torch.equal(idx, oi[:1024])
    kept = oi[1024:]
    assert torch.equal(R[kept], g[kept])  # withheld pairs stay in the residual


def test_segmented_per_tensor_mode():
    """threshold=0 mode: every tensor its own segment, one launch."""
    sizes = [64, 1000, 4097, 30000, 7, 2048]
    offs, total = [], 0
    for s in sizes:
        offs.append(total)
        total += (s + 63) // 64 * 64
    g = torch.zeros(total)
    r = torch.zeros(total)
    for i, (o, s) in enumerate(zip(offs, sizes)):
        g[o:o + s] = _data(s, 10 + i)
        r[o:o + s] = _data(s, 20 + i) * 0.5
    segs = [[(o, s)] for o, s in zip(offs, sizes)]
    for comp in ["gaussian", "topk"]:
        c, G, R, idx, val, _ = _run_cpu_chain(comp, g, r, 0.01, segments=segs, chunk=1024)
        exp_idx = []
        for o, s in zip(offs, sizes):
            acc, oi, ov, nr = oracle.compress(comp, g[o:o + s], r[o:o + s], 0.01)
            exp_idx.append(oi + o)
            assert torch.equal(R[o:o + s], nr)
        assert torch.equal(idx, torch.cat(exp_idx))


def test_merged_group_skips_padding():
    """A merged group spans several tensors; arena padding between them is not part of n."""
    sizes = [100, 5000, 33]
    offs = [0, 128, 5184]
    total = 5248
    g = torch.zeros(total)
    flat = _data(sum(sizes), 7)
    p = 0
    for o, s in zip(offs, sizes):
        g[o:o + s] = flat[p:p + s]
        p += s
    c, G, R, idx, val, _ = _run_cpu_chain("gaussian", g, torch.zeros(total), 0.01,
                                          segments=[[(o, s) for o, s in zip(offs, sizes)]], chunk=1024)
    oi = oracle.select("gaussian", flat, 0.01)
    # map dense group indices back to arena indices
    bounds = np.cumsum([0] + sizes)
    arena_idx = []
    for j in oi.tolist():
        t = int(np.searchsorted(bounds, j, side="right") - 1)
        arena_idx.append(offs[t] + j - bounds[t])
    assert idx.tolist() == arena_idx
    st = c.read_stats()[0]
    m, s = oracle.moments(flat)
    assert abs(st["mean"] - m) <= 1e-6 * max(1, abs(m)) and abs(st["std"] - s) <= 1e-6 * s


def test_scatter_mean_and_unscatter():
    n, P = 10000, 4
    G = torch.zeros(n)
    R = torch.zeros(n)
    c = ch.CompressChain(G, R, [[(0, n)]], compressor="topk", density=0.01)
    arena = ch.SlotBuffer(P, c.cap, "cpu")
    pairs = []
    for r in range(P):
        G.copy_(_data(n, 30 + r))
        R.zero_()
        c.ctrl.zero_()  # emulate P independent ranks that are all at sequence 0
        c.compress(ch.make_dst([arena.slot_ptr(r)], 0))
        i, v, _ = arena.read_slot(r)
        pairs.append((i, v))
    dense = torch.zeros(n)
    ch.scatter(arena.src(wait=True), c.ctrl, dense, 1.0 / P)
    assert torch.equal(dense, oracle.decompress_mean(pairs, n, P))
    ch.unscatter(arena.src(), c.ctrl, dense)
    assert torch.count_nonzero(dense) == 0


def test_randomk_distinct_and_ef():
    n = 50000
    g = _data(n, 40)
    G, R = g.clone(), torch.zeros(n)
    c = ch.CompressChain(G, R, [[(0, n)]], compressor="topk", density=0.01)
    sb = ch.SlotBuffer(1, c.cap, "cpu")
    c.randomk_send(ch.make_dst([sb.slot_ptr(0)], 0), seed=1234)
    idx, val, _ = sb.read_slot(0)
    assert idx.numel() == 500 and idx.unique().numel() == 500
    assert torch.equal(val, g[idx])
    assert torch.count_nonzero(R[idx]) == 0
    rebuilt = R.clone()
    rebuilt[idx] = val
    assert torch.equal(rebuilt, g)
    # a different seed draws a different set
    G.copy_(g); R.zero_()
    c.randomk_send(ch.make_dst([sb.slot_ptr(0)], 0), seed=99)
    idx2, _, _ = sb.read_slot(0)
    assert not torch.equal(idx.sort().values, idx2.sort().values)


def test_sgd_matches_torch_optim():
    torch.manual_seed(0)
    sizes = [(64, 3, 3, 3), (64,), (10, 64), (10,)]
    params = [torch.nn.Parameter(torch.randn(s)) for s in sizes]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in params]
    groups = [{"params": [ref[1], ref[3]], "weight_decay": 0.0}, {"params": [ref[0], ref[2]], "weight_decay": 1e-4}]
    opt = torch.optim.SGD(groups, lr=0.1, momentum=0.9)
    offs, total = [], 0
    for p in params:
        offs.append(total)
        total += (p.numel() + 63) // 64 * 64
    P, G, M = torch.zeros(total), torch.zeros(total), torch.zeros(total)
    for o, p in zip(offs, params):
        P[o:o + p.numel()] = p.detach().view(-1)
    tables = ch.FlatTables([(o, p.numel()) for o, p in zip(offs, params)], [1, 0, 1, 0], "cpu", chunk=512)
    hyp = [{"lr": 0.1, "momentum": 0.9, "weight_decay": 0.0}, {"lr": 0.1, "momentum": 0.9, "weight_decay": 1e-4}]
    for step in range(3):
        for o, p, q in zip(offs, params, ref):
            gr = torch.randn(p.shape)
            q.grad = gr.clone()
            G[o:o + p.numel()] = gr.view(-1)
        opt.step()
        ch.sgd_step(P, G, M, tables, ch.make_hyper(hyp, "cpu", first_step=(step == 0)))
        for o, q in zip(offs, ref):
            assert torch.allclose(P[o:o + q.numel()], q.detach().view(-1), rtol=0, atol=1e-7)


def test_lars_and_grad_norm():
    torch.manual_seed(1)
    sizes = [300, 4100, 17]
    offs = [0, 320, 4480]
    total = 4544
    P, G, A = torch.zeros(total), torch.zeros(total), torch.ones(total)
    ws, gs = [], []
    for o, s in zip(offs, sizes):
        w, g = torch.randn(s), torch.randn(s) * 0.1
        P[o:o + s], G[o:o + s] = w, g
        ws.append(w); gs.append(g)
    tables = ch.FlatTables(list(zip(offs, sizes)), [0, 0, 0], "cpu", chunk=1024)
    out = ch.grad_norm(G, tables, 0.25)
    tot = torch.cat(gs).norm(2)
    assert abs(out[0] - tot) < 1e-5 * tot and abs(out[1] - min(1.0, 0.25 / (tot + 1e-6))) < 1e-6
    hyp = [{"lr": 0.1, "momentum": 0.9, "weight_decay": 5e-4, "eeta": 0.001, "epsilon": 1e-5}]
    ch.lars_step(P, G, A, tables, ch.make_hyper(hyp, "cpu", True))
    for o, s, w, g in zip(offs, sizes, ws, gs):
        w2, a2 = oracle.lars_reference(w, g, torch.ones(s), 0.1, 0.9, 5e-4, 0.001, 1e-5)
        assert torch.allclose(P[o:o + s], w2, rtol=1e-6, atol=1e-7)
        assert torch.allclose(A[o:o + s], a2, rtol=1e-6, atol=1e-7)


def test_bucket_compressor():
    n = 5000
    g = _data(n, 50)
    G = g.clone()
    c = ch.CompressChain(G, torch.zeros(n), [[(0, n)]], compressor="topk", density=0.01)
    means = torch.zeros(2)
    c.bucket_means(means)
    assert torch.allclose(means[0], g[g >= 0].mean(), rtol=1e-6) and torch.allclose(means[1], g[g < 0].mean(), rtol=1e-6)
    gm = torch.tensor([0.02, -0.03])
    c.bucket_reconstruct(means, gm)
    exp = torch.where(g >= 0, (g - means[0]) + gm[0], (g - means[1]) + gm[1])
    assert torch.equal(G, exp)


# ----------------------------------------------------------- against the reference's own code
def test_oracle_matches_reference_gaussian(reference_modules):
    comp, utils = reference_modules
    for seed, kind, n, rho in [(1, "normal", 269722, 0.001), (2, "laplace", 100000, 0.004),
                               (3, "shifted", 50000, 0.001), (4, "normal", 1000, 0.001)]:
        g = _data(n, seed, kind)
        r = _data(n, seed + 100, kind) * 0.5
        comp.GaussianCompressor.clear()
        comp.GaussianCompressor.residuals["t"] = r.clone()
        t = g.clone()
        _, ridx, rval = comp.GaussianCompressor.compress(t, "t", ratio=rho)
        acc = g + r
        mean, std = float(torch.mean(acc)), float(torch.std(acc))  # the reference's own moments
        oi, _ = oracle.gaussian_select(acc, rho, 3, mean=mean, std=std)
        assert torch.equal(oi, ridx)
        assert torch.equal(acc[oi], rval)
        exp_res = acc.clone(); exp_res[oi] = 0
        assert torch.equal(comp.GaussianCompressor.residuals["t"], exp_res)
        # and our own double-precision moments agree to fp32 round-off
        m2, s2 = oracle.moments(acc)
        assert abs(m2 - mean) <= 2e-6 * max(abs(mean), std) and abs(s2 - std) <= 2e-6 * std


def test_oracle_matches_reference_topk_and_redsync(reference_modules):
    comp, utils = reference_modules
    g = _data(100000, 7)
    k = ch.compute_k(g.numel(), 0.001)
    rv, ri = comp.TopKCompressor.topk(g, k)  # the intended selector (compression.py:45-49)
    assert torch.equal(oracle.topk_select(g, k), ri.sort().values)
    for name, cls in [("redsync", comp.RedSyncCompressor), ("redsynctrim", comp.RedSyncTrimCompressor)]:
        cls.clear()
        comp.RedSyncCompressor.residuals = {}
        _, ridx, rval = cls.compress(g.clone(), "x", ratio=0.001)
        oi = oracle.select(name, g, 0.001)
        assert torch.equal(oi, ridx), name


def test_zabs_matches_reference(reference_modules):
    comp, utils = reference_modules
    for rho in [0.001, 0.004, 0.015625, 0.05, 0.25]:
        left, right = utils.gen_threshold_from_normal_distribution(1 - rho, 0.0, 1.0)
        assert right == ch.gaussian_zabs(rho)


# ---------------------------------------------------------------------------------------------- properties
from hypothesis import HealthCheck, given, settings as hsettings, strategies as st  # noqa: E402


@hsettings(max_examples=30, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(comp=st.sampled_from(COMPS + ["dgcsampling"]),
       n=st.integers(min_value=1, max_value=40000),
       density=st.sampled_from([0.001, 0.004, 0.015625, 0.05]),
       kind=st.sampled_from(["normal", "laplace", "shifted", "ties"]),
       chunk=st.sampled_from([128, 1024, 8192]),
       seed=st.integers(min_value=0, max_value=2 ** 16))
def test_compress_invariants_hold_for_any_input(comp, n, density, kind, chunk, seed):
    """Whatever the selector, size, density, data and block table: what was sent plus what stayed equals gradient plus
    old residual bit for bit, indices come out unique and ascending, the count respects the slot capacity and the
    header, the gradient arena is consumed, and exact top-k sends exactly k."""
    g, r = _data(n, seed, kind), _data(n, seed + 1, kind) * 0.5
    c, G, R, idx, val, seq = _run_cpu_chain(comp, g, r, density, chunk=chunk)
    ctrl = c.read_ctrl()
    acc = g + r
    assert seq == 1 and ctrl["sent"] == idx.numel() <= c.cap and ctrl["total"] >= ctrl["sent"]
    if idx.numel():
        assert int(idx.min()) >= 0 and int(idx.max()) < n
        assert torch.all(idx[1:] > idx[:-1])
        assert torch.equal(val, acc[idx])                      # values travel unchanged
    rebuilt = R.clone()
    rebuilt[idx] += val
    assert torch.equal(rebuilt, acc)                           # nothing lost, nothing invented
    assert torch.count_nonzero(R[idx]) == 0                    # sent coordinates leave the residual
    assert torch.count_nonzero(G) == 0
    if comp == "topk" and ctrl["total"] <= c.cap:
        k = ch.compute_k(n, density)
        assert idx.numel() == min(k, n)
        if 0 < idx.numel() < n:
            rest = torch.ones(n, dtype=torch.bool)
            rest[idx] = False
            assert acc.abs()[idx].min() >= acc.abs()[rest].max()


def _fill_cpu_arena(arena, lists, seq):
    from gaussiank_sgd_b200.ops import native as N
    P = len(lists)
    for r, (sel, v) in enumerate(lists):
        words = sel.numpy().astype(np.uint64) | (v.numpy().view(np.uint32).astype(np.uint64) << np.uint64(32))
        base = ((seq & 1) * P + r) * arena.slot_stride
        arena.t[base] = int(sel.numel()) | (seq << 32)
        arena.t[base + N.GK_SLOT_HDR: base + N.GK_SLOT_HDR + sel.numel()] = torch.from_numpy(words.view(np.int64))


@hsettings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(P=st.integers(min_value=1, max_value=8), k=st.integers(min_value=0, max_value=400),
       momentum=st.sampled_from([0.0, 0.9]), nesterov=st.booleans(), wd=st.sampled_from([0.0, 1e-3]),
       first=st.booleans(), seed=st.integers(min_value=0, max_value=2 ** 16))
def test_folded_update_equals_scatter_then_sgd_for_any_lists(P, k, momentum, nesterov, wd, first, seed):
    """gkcpu_sgd_recv (receive folded into the update, no dense gradient) against gkcpu_scatter + gkcpu_sgd for random
    rank counts, list lengths (empty lists and coordinates held by several ranks included) and hyper-parameters:
    coordinates nobody sent are bit-identical, the others agree to rounding, the mask arena comes back zero."""
    if nesterov and momentum == 0.0:
        nesterov = False
    g = torch.Generator().manual_seed(seed)
    sizes = [130, 4097, 17, 2500]
    offs, total = [], 0
    for s in sizes:
        offs.append(total)
        total += (s + 63) // 64 * 64
    tables = ch.FlatTables(list(zip(offs, sizes)), [0, 1, 0, 1], "cpu")
    hyp = [{"lr": 0.05, "momentum": momentum, "weight_decay": 0.0, "nesterov": nesterov},
           {"lr": 0.1, "momentum": momentum, "weight_decay": wd, "nesterov": nesterov}]
    valid = torch.cat([torch.arange(o, o + s) for o, s in zip(offs, sizes)])
    shared = valid[torch.randperm(valid.numel(), generator=g)[:min(k, 50)]]
    lists = []
    for r in range(P):
        sel = valid[torch.randperm(valid.numel(), generator=g)[:k]]
        if r % 2:
            sel = torch.cat([sel, shared])
        sel = sel.unique()
        lists.append((sel, torch.randn(sel.numel(), generator=g)))
    cap = max([l[0].numel() for l in lists] + [1]) + 8
    arena = ch.SlotBuffer(P, cap, "cpu", parities=2)
    ctrl = torch.zeros(ch.CTRL_DTYPE.itemsize, dtype=torch.uint8)
    seq = 5
    ctrl.view(torch.int32)[10] = seq
    _fill_cpu_arena(arena, lists, seq)
    hy = ch.make_hyper(hyp, "cpu", first)
    P0, M0 = torch.randn(total, generator=g), torch.randn(total, generator=g) * 0.1
    mom = (lambda M: M) if momentum else (lambda M: None)
    P1, M1, G = P0.clone(), M0.clone(), torch.zeros(total)
    ch.scatter(arena.src(wait=True, timeout_s=2.0), ctrl, G, 1.0 / P)
    ch.sgd_step(P1, G, mom(M1), tables, hy)
    P2, M2, mask = P0.clone(), M0.clone(), torch.zeros(total)
    ch.sgd_recv_step(P2, mask.view(torch.int32), mom(M2), tables, hy, arena.src(wait=True, timeout_s=2.0), ctrl, 1.0 / P)
    assert torch.count_nonzero(mask) == 0 and int(ctrl.view(torch.int32)[11]) == 0
    touched = torch.zeros(total, dtype=torch.bool)
    for sel, _ in lists:
        touched[sel] = True
    assert torch.equal(P1[~touched], P2[~touched]) and torch.equal(M1[~touched], M2[~touched])
    eps = 4 * torch.finfo(torch.float32).eps
    assert (P1 - P2).abs().max() <= eps * max(1.0, float(P1.abs().max()))
    assert (M1 - M2).abs().max() <= eps * max(1.0, float(M1.abs().max()))
