#!/usr/bin/env python
"""Launch every kernel of libgk_cuda.so's chain / optimizer families a few times on ResNet-50-sized arenas, for

    ncu --set full --clock-control none --import-source on -k regex:k_ -o gpurun_out/prof_all python tools/ncu_all.py

(the fused BN kernels are captured from a training step: ``tools/profile_step.py --no-graph`` under
``ncu -k regex:k_bn|k_pool``).  Run it once without ncu first; it prints what it launched.
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200.ops import chain as ch  # noqa: E402
from gaussiank_sgd_b200.ops import native as N  # noqa: E402

REPS = 2


def main():
    dev = "cuda"
    n = 25557032
    total = (n + 63) // 64 * 64
    torch.manual_seed(0)
    G0 = torch.randn(total, device=dev) * 0.01
    R0 = torch.randn(total, device=dev) * 0.003
    P = torch.randn(total, device=dev)
    M = torch.zeros(total, device=dev)
    tables = ch.FlatTables([(0, n)], [0], dev)
    hyper = ch.make_hyper([{"lr": 0.01, "momentum": 0.875, "weight_decay": 6.1e-5, "eeta": 0.001, "epsilon": 1e-5}],
                          dev, False).to(dev)
    st = torch.cuda.current_stream().cuda_stream
    lib = N.cuda_lib()
    done = []

    def chain(comp, density=0.001):
        G, R = G0.clone(), R0.clone()
        c = ch.CompressChain(G, R, [[(0, n)]], compressor=comp, density=density)
        sb = ch.SlotBuffer(1, c.cap, dev, parities=2)
        return c, G, R, sb, ch.make_dst([sb.slot_ptr(0)], sb.parity_stride)

    for comp in ("gaussian", "topk", "dgcsampling", "redsync"):
        c, G, R, sb, dst = chain(comp)
        for _ in range(REPS):
            G.copy_(G0); R.copy_(R0)
            c.compress(dst, 7)
        done.append(f"compress[{comp}]")
        if comp == "gaussian":
            dense = torch.zeros(total, device=dev)
            for _ in range(REPS):
                ch.scatter(sb.src(), c.ctrl, dense, 1.0)
                ch.unscatter(sb.src(), c.ctrl, dense)
                ch.sgd_recv_step(P, dense.view(torch.int32), M, tables, hyper, sb.src(), c.ctrl, 1.0)
            done.append("scatter / unscatter / sgd_recv")
            # rank-ordered receive (one launch per source rank) on an emulated 4-rank arena
            arena = ch.SlotBuffer(4, c.cap, dev, parities=2)
            seq = c.read_ctrl()["seq"]
            for r in range(4):
                c.ctrl.view(torch.int32)[10] = seq
                G.copy_(G0 * (1 + 0.1 * r)); R.copy_(R0)
                c.compress(ch.make_dst([arena.slot_ptr(r)], arena.parity_stride))
            for _ in range(REPS):
                ch.scatter(arena.src(wait=True, timeout_s=2.0), c.ctrl, dense, 0.25)
                ch.unscatter(arena.src(), c.ctrl, dense)
            done.append("scatter ordered x4")
        del c, G, R, sb
    # random-k
    c, G, R, sb, dst = chain("threshold")
    for _ in range(REPS):
        G.copy_(G0)
        c.randomk_send(dst, 11)
    done.append("randomk")
    # bucket compressor
    means, gm = torch.zeros(2, device=dev), torch.tensor([0.01, -0.01], device=dev)
    for _ in range(REPS):
        c.bucket_means(means)
        c.bucket_reconstruct(means, gm)
    done.append("bucket")
    # optimizer kernels
    A = torch.ones(total, device=dev)
    for _ in range(REPS):
        ch.sgd_step(P, G0, M, tables, hyper)
        ch.lars_step(P, G0, A, tables, hyper)
        out = ch.grad_norm(G0, tables, 0.25)
        ch.scale_(G, tables, out[1:2])
    done.append("sgd / lars / grad_norm / scale")
    hist = torch.zeros(256, dtype=torch.int32, device=dev)
    for _ in range(REPS):
        N.check(lib.gk_histogram(R0.data_ptr(), n, C.c_float(-0.05), C.c_float(0.05), 256, hist.data_ptr(), st))
        N.check(lib.gk_copy(G.data_ptr(), G0.data_ptr(), n, st))
    done.append("histogram / copy")
    torch.cuda.synchronize()
    print("launched:", "; ".join(done))


if __name__ == "__main__":
    main()
