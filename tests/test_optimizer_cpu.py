"""DistributedOptimizer on the CPU: single process with force_compress against a hand-rolled oracle
loop, and two gloo workers (BASELINE.json config 1 in miniature)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp
import torch.nn as nn

import gaussiank_sgd_b200.distributed_optimizer as hvd
from gaussiank_sgd_b200.compression import compressors
from gaussiank_sgd_b200.ops import oracle


def _mlp(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(20, 64), nn.ReLU(), nn.Linear(64, 64), nn.ReLU(), nn.Linear(64, 5))


def _param_groups(model, wd):
    decay = [p for n, p in model.named_parameters() if p.dim() > 1]
    no_decay = [p for n, p in model.named_parameters() if p.dim() <= 1]
    return [{"params": no_decay, "weight_decay": 0.0}, {"params": decay, "weight_decay": wd}]


def _oracle_step(model, opt, resid, comp, density, P=1):
    """Reference semantics with plain torch ops on the merged buffer (reverse parameter order)."""
    params = [p for _, p in model.named_parameters()][::-1]
    g = torch.cat([p.grad.reshape(-1) for p in params])
    acc, idx, vals, new_resid = oracle.compress(comp, g, resid, density)
    dense = oracle.decompress_mean([(idx, vals)] * 1, g.numel(), P)
    off = 0
    for p in params:
        p.grad = dense[off:off + p.numel()].view_as(p).clone()
        off += p.numel()
    opt.step()
    return new_resid, idx.numel()


@pytest.mark.parametrize("comp", ["gaussian", "topk", "redsync"])
@pytest.mark.parametrize("threshold", [524288000, 0])
def test_single_process_matches_oracle(comp, threshold):
    density = 0.05
    m1, m2 = _mlp(), _mlp()
    o1 = torch.optim.SGD(_param_groups(m1, 1e-4), lr=0.1, momentum=0.9)
    o2 = torch.optim.SGD(_param_groups(m2, 1e-4), lr=0.1, momentum=0.9)
    opt = hvd.DistributedOptimizer(o1, named_parameters=m1.named_parameters(), compression=compressors[comp],
                                   is_sparse=True, density=density, threshold=threshold,
                                   dynamic_densities=None, force_compress=True)
    n = sum(p.numel() for p in m2.parameters())
    resid = torch.zeros(n)
    x = torch.randn(16, 20)
    y = torch.randint(0, 5, (16,))
    lossf = nn.CrossEntropyLoss()
    for step in range(4):
        opt.zero_grad()
        lossf(m1(x), y).backward()
        opt.step()
        if threshold != 0:
            o2.zero_grad()
            lossf(m2(x), y).backward()
            resid, cnt = _oracle_step(m2, o2, resid, comp, density)
            for (n1, p1), (n2, p2) in zip(m1.named_parameters(), m2.named_parameters()):
                assert torch.allclose(p1, p2, rtol=0, atol=1e-6), (step, n1)
    sel = opt.selected_num_gradients()
    assert len(sel) == 4 and all(s > 0 for s in sel)
    opt.close()


def test_plain_optimizer_when_single_worker():
    """size()==1 and no force_compress: the wrapper is a plain (fused) SGD (distributed_optimizer.py:97)."""
    m1, m2 = _mlp(), _mlp()
    o1 = torch.optim.SGD(_param_groups(m1, 1e-4), lr=0.05, momentum=0.9)
    o2 = torch.optim.SGD(_param_groups(m2, 1e-4), lr=0.05, momentum=0.9)
    opt = hvd.DistributedOptimizer(o1, named_parameters=m1.named_parameters(), compression=compressors["gaussian"],
                                   is_sparse=True, density=0.001)
    x, y = torch.randn(8, 20), torch.randint(0, 5, (8,))
    for _ in range(3):
        for o, m in ((opt, m1), (o2, m2)):
            o.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    for p1, p2 in zip(m1.parameters(), m2.parameters()):
        assert torch.allclose(p1, p2, rtol=0, atol=1e-6)
    sd = opt.state_dict()
    assert len(sd["state"]) == 6 and "momentum_buffer" in sd["state"][0]


def test_gradient_accumulation_local_flag():
    m1 = _mlp()
    opt = hvd.DistributedOptimizer(torch.optim.SGD(m1.parameters(), lr=0.1), named_parameters=m1.named_parameters(),
                                   compression=compressors["topk"], is_sparse=True, density=0.1,
                                   dynamic_densities=None, force_compress=True)
    x, y = torch.randn(8, 20), torch.randint(0, 5, (8,))
    opt.zero_grad()
    opt.local = True
    nn.functional.cross_entropy(m1(x), y).backward()
    g1 = torch.cat([p.grad.reshape(-1).clone() for p in m1.parameters()])
    opt.local = False
    nn.functional.cross_entropy(m1(x), y).backward()
    g2 = torch.cat([p.grad.reshape(-1).clone() for p in m1.parameters()])
    assert torch.allclose(g2, 2 * g1, rtol=1e-5, atol=1e-7)   # accumulated, nothing compressed yet
    opt.synchronize()
    gs = torch.cat([p.grad.reshape(-1) for p in m1.parameters()])
    k = max(int(gs.numel() * 0.1), 1)
    assert int((gs != 0).sum()) == k
    opt.close()


def test_density_schedule():
    m = _mlp()
    opt = hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.1), named_parameters=m.named_parameters(),
                                   compression=compressors["gaussian"], is_sparse=True, density=0.001,
                                   force_compress=True)
    assert opt.get_current_density() == 0.015625
    opt.increase_one_epoch()
    assert opt.get_current_density() == 0.004
    opt.increase_one_epoch(); opt.increase_one_epoch()
    assert opt.get_current_density() == 0.001
    opt.close()


# ------------------------------------------------------------------ two gloo workers
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, comp, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), GK_FORCE_CPU="1")
    torch.set_num_threads(1)
    hvd.init()
    model = _mlp(seed=rank)            # different initialisations on purpose
    base = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
    is_sparse = comp != "none"
    opt = hvd.DistributedOptimizer(base, named_parameters=model.named_parameters(), compression=compressors[comp],
                                   is_sparse=is_sparse, density=0.05 if is_sparse else 1, threshold=524288000,
                                   dynamic_densities=None)
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    torch.manual_seed(100 + rank)
    x, y = torch.randn(8, 20), torch.randint(0, 5, (8,))
    grads_after_sync = None
    for step in range(3):
        opt.zero_grad()
        nn.functional.cross_entropy(model(x), y).backward()
        raw = torch.cat([p.grad.reshape(-1).clone() for p in list(model.parameters())[::-1]])
        opt.synchronize()
        grads_after_sync = torch.cat([p.grad.reshape(-1).clone() for p in list(model.parameters())[::-1]])
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    q.put((rank, flat.numpy().copy(), raw.numpy().copy(), grads_after_sync.numpy().copy()))
    hvd.barrier()
    opt.close()
    hvd.shutdown()


@pytest.mark.parametrize("comp", ["gaussian", "topk", "none"])
def test_two_gloo_workers_stay_in_sync(comp):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, comp, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, w0, raw0, g0), (_, w1, raw1, g1) = [(r, *map(torch.from_numpy, t)) for r, *t in res]
    assert torch.equal(w0, w1)            # replicas stay bit-identical
    assert torch.equal(g0, g1)            # both ranks decompress the same averaged gradient
    if comp == "none":
        # last step: averaged dense gradient
        assert torch.allclose(g0, (raw0 + raw1) / 2, rtol=1e-6, atol=1e-8)
    else:
        nz = int((g0 != 0).sum())
        assert 0 < nz <= 2 * max(int(g0.numel() * 0.05 * 4 / 3) + 1, 1) * 2


# ------------------------------------------------------------------ round 2
def test_folded_step_equals_materialised_step_cpu():
    """step() folds the receive into the update (no dense gradient); synchronize()+step() materialises it.
    Same parameters up to the rounding of the split FMA, and p.grad stays zero on the folded path."""
    res = {}
    for fold in (True, False):
        m = _mlp()
        o = torch.optim.SGD(_param_groups(m, 1e-3), lr=0.1, momentum=0.9, nesterov=True)
        opt = hvd.DistributedOptimizer(o, named_parameters=m.named_parameters(), compression=compressors["topk"],
                                       is_sparse=True, density=0.05, threshold=524288000, dynamic_densities=None,
                                       force_compress=True)
        torch.manual_seed(1)
        x, y = torch.randn(16, 20), torch.randint(0, 5, (16,))
        for _ in range(4):
            opt.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
            if fold:
                assert opt.can_fold()
            else:
                opt.synchronize()
                assert not opt.can_fold()
            opt.step()
            if fold:
                assert all(int(torch.count_nonzero(p.grad)) == 0 for p in m.parameters())
        res[fold] = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
        opt.close()
    assert torch.allclose(res[True], res[False], rtol=0, atol=1e-6)
    assert not torch.equal(res[True], torch.cat([p.detach().reshape(-1) for p in _mlp().parameters()]))


@pytest.mark.parametrize("comp", ["uniform_abs_topk", "bucketized_topk"])
def test_ordering_selectors_run_their_own_algorithm(comp):
    """No registered key may silently run another algorithm: the fork's ordering selectors go through the
    registry object's compress() (torch ops) inside the optimizer and select what that function selects."""
    if comp not in compressors:
        pytest.skip("selector not registered")
    m = _mlp()
    o = torch.optim.SGD(m.parameters(), lr=0.1)
    opt = hvd.DistributedOptimizer(o, named_parameters=m.named_parameters(), compression=compressors[comp],
                                   is_sparse=True, density=0.05, threshold=524288000, dynamic_densities=None,
                                   force_compress=True)
    assert opt._torch_select and not opt.graph_safe
    x, y = torch.randn(16, 20), torch.randint(0, 5, (16,))
    compressors[comp].clear()
    opt.zero_grad()
    nn.functional.cross_entropy(m(x), y).backward()
    names = opt._groups[0].names
    raw = torch.cat([dict(m.named_parameters())[n].grad.reshape(-1).clone() for n in names])
    opt.synchronize()
    got = torch.cat([dict(m.named_parameters())[n].grad.reshape(-1) for n in names])
    compressors[comp].clear()
    _t, idx, vals = compressors[comp].compress(raw.clone(), "check", ratio=0.05)
    exp = torch.zeros_like(raw)
    exp[idx.reshape(-1)] = vals.reshape(-1)
    assert torch.equal(got, exp)
    exact = oracle.topk_select(raw, max(int(raw.numel() * 0.05), 1))
    assert not torch.equal(torch.sort(idx.reshape(-1)).values, exact)      # really not exact top-k
    compressors[comp].clear()
    opt.close()


def test_unknown_compressor_is_refused():
    m = _mlp()

    class Fake:
        key = "does_not_exist"

    with pytest.raises(ValueError):
        hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.1), named_parameters=m.named_parameters(),
                                 compression=Fake(), is_sparse=True, density=0.05, force_compress=True)


def _resume_worker(rank, world, port, ckpt, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), GK_FORCE_CPU="1")
    torch.set_num_threads(1)
    hvd.init()

    def make():
        model = _mlp(seed=3)
        base = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
        opt = hvd.DistributedOptimizer(base, named_parameters=model.named_parameters(),
                                       compression=compressors["topk"], is_sparse=True, density=0.05,
                                       threshold=524288000, dynamic_densities=None)
        return model, opt

    def steps(model, opt, n, seed):
        g = torch.Generator().manual_seed(seed + rank)
        for _ in range(n):
            x, y = torch.randn(8, 20, generator=g), torch.randint(0, 5, (8,), generator=g)
            opt.zero_grad()
            nn.functional.cross_entropy(model(x), y).backward()
            opt.step()

    model, opt = make()
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    steps(model, opt, 3, 10)
    if rank == 0:
        torch.save({"state": model.state_dict(), "optimizer": opt.state_dict(), "gk_extra": opt.extra_state_dict()}, ckpt)
    hvd.barrier()
    opt.close()
    # ---- resume the way ssgd() does: rank 0 alone reads the file
    model, opt = make()
    if rank == 0:
        ck = torch.load(ckpt, weights_only=False)
        model.load_state_dict(ck["state"])
        opt.load_state_dict(ck["optimizer"])
        opt._init_fused_state_from_loaded()
        opt.load_extra_state_dict(ck["gk_extra"])
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    opt.sync_resume_state(root_rank=0)
    assert opt._steps_done == 3
    steps(model, opt, 3, 20)
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    mom = opt._mom.clone()
    q.put((rank, flat.numpy().copy(), mom.numpy().copy()))
    hvd.barrier()
    opt.close()
    hvd.shutdown()


def test_resume_with_two_workers_keeps_replicas_identical(tmp_path):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ckpt = str(tmp_path / "resume.pth")
    procs = [ctx.Process(target=_resume_worker, args=(r, 2, port, ckpt, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert (res[0][1] == res[1][1]).all() and (res[0][2] == res[1][2]).all()
    assert abs(res[0][2]).sum() > 0


def test_dense_synchronize_is_idempotent():
    """synchronize(); clip; step() calls synchronize() twice per iteration (dist_trainer.py:80-86): the second
    call must not exchange again (bucket: subtract/add the means twice; none: a second all-reduce)."""
    m = _mlp()
    opt = hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.1), named_parameters=m.named_parameters(),
                                   compression=compressors["bucket"], is_sparse=False, density=1,
                                   dynamic_densities=None, force_compress=True)
    x, y = torch.randn(8, 20), torch.randint(0, 5, (8,))
    opt.zero_grad()
    nn.functional.cross_entropy(m(x), y).backward()
    opt.synchronize()
    g1 = torch.cat([p.grad.reshape(-1).clone() for p in m.parameters()])
    it = opt.train_iter
    opt.synchronize()
    g2 = torch.cat([p.grad.reshape(-1).clone() for p in m.parameters()])
    assert torch.equal(g1, g2) and opt.train_iter == it
    opt.step()
    opt.close()


def test_verify_mode_cross_checks_against_the_oracle(monkeypatch):
    """GK_VERIFY_EVERY: every N-th step materialises the gradient and compares selection, received gradient
    and (with >1 rank) replica checksums with the oracle."""
    from gaussiank_sgd_b200 import settings
    monkeypatch.setattr(settings, "VERIFY_EVERY", 2)
    m = _mlp()
    opt = hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9),
                                   named_parameters=m.named_parameters(), compression=compressors["gaussian"],
                                   is_sparse=True, density=0.05, threshold=524288000, dynamic_densities=None,
                                   force_compress=True)
    x, y = torch.randn(16, 20), torch.randint(0, 5, (16,))
    for _ in range(5):
        opt.zero_grad()
        nn.functional.cross_entropy(m(x), y).backward()
        opt.step()
    assert opt.verified_steps == 3          # iterations 0, 2, 4
    opt.close()


def test_receive_arena_follows_the_density_schedule():
    """Slot capacity is 8x k of the CURRENT density: after the warm-up epochs the arena shrinks with it, the
    graph version is bumped (captured steps must be re-captured) and training simply continues."""
    torch.manual_seed(0)
    m = nn.Sequential(nn.Linear(64, 512), nn.ReLU(), nn.Linear(512, 512), nn.ReLU(), nn.Linear(512, 10))
    opt = hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9),
                                   named_parameters=m.named_parameters(), compression=compressors["topk"],
                                   is_sparse=True, density=0.001, threshold=524288000, force_compress=True)
    n = sum(p.numel() for p in m.parameters())
    x, y = torch.randn(8, 64), torch.randint(0, 10, (8,))
    caps, versions = [], []
    for epoch in range(3):
        for _ in range(2):
            opt.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
            opt.step()
        caps.append(opt._groups[0].chain.cap)
        versions.append(opt.graph_version)
        assert opt._groups[0].transport.src().slot_stride == caps[-1] + 2
        opt.increase_one_epoch()
    assert caps == [max(8 * int(n * d), 1024) for d in (0.015625, 0.004, 0.001)]
    assert versions == [0, 1, 2]
    assert opt.selected_num_gradients()[-1] == max(int(n * 0.001), 1)
    assert opt.read_overflow() == 0
    opt.close()
