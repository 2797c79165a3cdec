#!/usr/bin/env python
"""Per-kernel timing of the sparsification chain and the fused optimizer on one GPU (CUDA events, L2
flushed between iterations).  Prints achieved GB/s against MEASURED_PEAKS.json.

    python tools/bench_chain.py [--n 25557032] [--comp gaussian] [--density 0.001] [--chunk 8192]
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200.ops import chain as ch  # noqa: E402
from gaussiank_sgd_b200.ops import native as N  # noqa: E402
from gaussiank_sgd_b200.utils.perf_model import measured_peaks  # noqa: E402


def timed(fn, flush, iters=10, prep=None):
    ts = []
    for _ in range(iters):
        if prep is not None:
            prep()
        if flush is not None:
            flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=25557032)
    ap.add_argument("--comp", default="gaussian")
    ap.add_argument("--density", type=float, default=0.001)
    ap.add_argument("--chunk", type=int, default=ch.DEFAULT_CHUNK)
    ap.add_argument("--segments", type=int, default=1)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--json", default=None)
    ap.add_argument("--timeline", action="store_true")
    a = ap.parse_args()
    dev = "cuda"
    n = a.n
    total = (n + 63) // 64 * 64
    G = torch.randn(total, device=dev) * 0.01
    R = torch.randn(total, device=dev) * 0.003
    P = torch.randn(total, device=dev)
    M = torch.zeros(total, device=dev)
    if a.segments == 1:
        segs = [[(0, n)]]
    else:
        per = (n // a.segments) // 64 * 64
        segs = [[(i * per, per)] for i in range(a.segments)]
    c = ch.CompressChain(G, R, segs, compressor=a.comp, density=a.density, chunk=a.chunk)
    sb = ch.SlotBuffer(1, c.cap, dev, parities=2)
    dst = ch.make_dst([sb.slot_ptr(0)], sb.parity_stride)
    tables = ch.FlatTables([(0, n)], [0], dev, chunk=a.chunk)
    hyper = ch.make_hyper([{"lr": 0.01, "momentum": 0.875, "weight_decay": 6.1e-5}], dev, False).to(dev)
    flush = None if a.no_flush else torch.zeros(64 * 1024 * 1024, device=dev)   # 256 MB > 126 MB L2
    peaks = measured_peaks()
    hbm = peaks["hbm_gbs"]
    rows = []

    G0, R0 = G.clone(), R.clone()

    def restore():          # every timed call sees the same fresh gradient + residual
        G.copy_(G0)
        R.copy_(R0)

    def report(name, fn, bytes_per_elem, prep=None):
        if prep:
            prep()
        fn(); torch.cuda.synchronize()
        med, best = timed(fn, flush, prep=prep)
        gbs = bytes_per_elem * n / (med * 1e-3) / 1e9
        rows.append({"kernel": name, "ms_median": round(med, 4), "ms_best": round(best, 4), "bytes_per_elem": bytes_per_elem,
                     "gbs": round(gbs, 1), "frac_of_%s_hbm" % peaks["source"]: round(gbs / hbm, 3)})
        print("%-28s %8.4f ms (best %8.4f)  %5.1f B/elem  %7.1f GB/s  %.3f of %s HBM" % (
            name, med, best, bytes_per_elem, gbs, gbs / hbm, peaks["source"]))

    topk = a.comp in ("topk", "mink")
    dgc = a.comp == "dgcsampling"
    def prep_k2():
        restore(); c.accumulate()

    def prep_k3():
        restore(); c.accumulate()
        if dgc:
            N.check(c._lib.gk_dgc_select(C.byref(c._desc), C.c_uint64(1), torch.cuda.current_stream().cuda_stream))
        elif topk:
            c.topk_select()
        else:
            c.count_pick()

    report("K1 accumulate+moments", c.accumulate, 16.0, prep=restore)
    if dgc:
        report("K9 dgc sample + select", lambda: N.check(c._lib.gk_dgc_select(
            C.byref(c._desc), C.c_uint64(1), torch.cuda.current_stream().cuda_stream)), 12.0, prep=prep_k2)
    elif topk:
        report("K4 radix x2 + count(gt,eq)", c.topk_select, 12.0, prep=prep_k2)
    else:
        report("K2 count+pick", c.count_pick, 4.0, prep=prep_k2)
    if a.timeline:
        # globaltimer stamps written by the kernels when GkChain.rsv has bit 2 set (debug builds of the numbers only)
        c._desc.rsv = 4
        for name, fn, prep, idx in (("K2", c.count_pick, prep_k2, (44, 58, 61, 62, 63, 59, 60, 45, 46, 47, 48, 49)),
                                    ("K3", lambda: c.compact_send(dst), prep_k3, (50, 51, 52))):
            prep()
            for i in idx:
                c.ctrl.view(torch.int32)[18 + i] = 0
            flush.add_(1.0) if flush is not None else None
            torch.cuda.synchronize()
            fn()
            torch.cuda.synchronize()
            ring = c.read_ctrl()["ring"]
            t0 = ring[idx[0]]
            print(name, "timeline ns since first block start:", [(ring[i] - t0) & 0xffffffff for i in idx[1:]])
        c._desc.rsv = 0
    report("K3 compact+send", lambda: c.compact_send(dst), 4.0, prep=prep_k3)
    report("chain (K1..K3)", lambda: c.compress(dst, 1), 32.0 if (topk or dgc) else 24.0, prep=restore)
    dense = torch.zeros(total, device=dev)
    report("K5 scatter (recv)", lambda: ch.scatter(sb.src(), c.ctrl, dense, 1.0), 0.0)
    report("unscatter (zero_grad)", lambda: ch.unscatter(sb.src(), c.ctrl, dense), 0.0)
    report("K6 fused SGD", lambda: ch.sgd_step(P, G, M, tables, hyper), 20.0)
    dense.zero_()
    report("K6 folded recv+SGD", lambda: ch.sgd_recv_step(P, dense.view(torch.int32), M, tables, hyper, sb.src(), c.ctrl,
                                                          1.0), 16.0)
    assert int(torch.count_nonzero(dense)) == 0
    if a.timeline:
        for i in range(53, 58):
            c.ctrl.view(torch.int32)[18 + i] = 0
        flush.add_(1.0) if flush is not None else None
        torch.cuda.synchronize()
        u = N.GkRecvUpd(P.data_ptr(), M.data_ptr(), dense.data_ptr(), tables.group_blk_t.data_ptr(), hyper.data_ptr(),
                        tables.gmap.data_ptr(), c.ctrl.data_ptr(), tables.bar.data_ptr(), 1.0, tables.nblk, dense.numel(), 4)
        src = sb.src()
        N.check(N.cuda_lib().gk_sgd_recv(C.byref(u), C.byref(src), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        ring = c.read_ctrl()["ring"]
        print("K6 folded timeline ns since first block start (phase A end, pass 1 end, barrier exit, end):",
              [(ring[i] - ring[53]) & 0xffffffff for i in range(54, 58)])
    report("K8 grad norm", lambda: ch.grad_norm(G, tables, 1.0), 4.0)
    report("memset (torch zero_)", lambda: G.zero_(), 4.0)
    report("copy (torch, ref)", lambda: M.copy_(P), 8.0)
    print("ctrl:", {k: v for k, v in c.read_ctrl().items() if k in ("total", "sent", "seq", "err")}, "k =", c.k_total)
    if a.json:
        with open(a.json, "w") as f:
            json.dump({"n": n, "comp": a.comp, "density": a.density, "chunk": a.chunk, "segments": a.segments,
                       "l2_flush": not a.no_flush, "peaks": peaks, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
