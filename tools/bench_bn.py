#!/usr/bin/env python
"""Per-shape timing of the fused BN(+residual)+ReLU kernels (csrc/gk_bn.cu) against the PyTorch/cuDNN
sequence they replace, on the activation shapes of ResNet-50 at batch 32.  CUDA events, L2 flushed
between iterations, roofline = measured copy bandwidth (MEASURED_PEAKS.json or the recipe's fallback).

    python tools/bench_bn.py [--batch 32] [--iters 20] [--out profiles/bn_fused_vs_cudnn.json]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200 import settings  # noqa: E402
from gaussiank_sgd_b200.ops.bn import bn_act, phase_stamps  # noqa: E402

SHAPES = [  # C, HW, residual, count in ResNet-50
    (64, 112, False, 1), (64, 56, False, 6), (256, 56, False, 1), (256, 56, True, 3),
    (128, 56, False, 1), (128, 28, False, 7), (512, 28, False, 1), (512, 28, True, 4),
    (256, 28, False, 1), (256, 14, False, 11), (1024, 14, False, 1), (1024, 14, True, 6),
    (512, 14, False, 1), (512, 7, False, 5), (2048, 7, False, 1), (2048, 7, True, 3)]


def peak_bw():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        d = json.load(open(p))
        for k in ("hbm_gbs", "copy_bandwidth_GBps", "hbm_copy_GBps", "copy_GBps"):
            if k in d:
                return float(d[k]) * 1e9
    except Exception:
        pass
    return 6.2e12     # recipe fallback: measured D2D copy, read + write bytes


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default=None)
    ap.add_argument("--once", action="store_true", help="one profiled forward+backward per shape (for ncu)")
    ap.add_argument("--qcap", type=int, default=None, help="geometry knob: log2 of the widest channel slab (quads)")
    ap.add_argument("--private-kb", type=int, default=None, help="geometry knob: barrier-free path up to this slab")
    ap.add_argument("--big-mb", type=int, default=None, help="geometry knob: 1024 threads per block from this many MB")
    ap.add_argument("--quiet", action="store_true", help="print one line per shape: graph ms fused / torch")
    a = ap.parse_args()
    if a.qcap is not None or a.private_kb is not None:
        from gaussiank_sgd_b200.ops import native
        native.check(native.cuda_lib().gk_bn_set_tuning(9 if a.qcap is None else a.qcap,
                                                        128 if a.private_kb is None else a.private_kb))
    if a.big_mb is not None:
        from gaussiank_sgd_b200.ops import native
        native.check(native.cuda_lib().gk_bn_set_big_mb(a.big_mb))
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev).view(torch.float32)
    bw = peak_bw()
    rows, tot = [], {"fused": 0.0, "torch": 0.0}
    for C, HW, res, cnt in SHAPES:
        x = torch.randn(a.batch, C, HW, HW, device=dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        r = torch.randn_like(x).requires_grad_(True) if res else None
        dy = torch.randn_like(x)
        bn = nn.BatchNorm2d(C).to(dev)
        out = {}

        def fused_f():
            out["y"] = bn_act(bn, x, residual=r, relu=True)

        def fused_b():
            torch.autograd.grad(out["y"], [x, bn.weight, bn.bias] + ([r] if res else []), dy, retain_graph=True)

        def torch_f():
            y = bn(x)
            if res:
                y = y + r
            out["y"] = F.relu(y, inplace=True)

        if a.once:          # for ncu --profile-from-start off: exactly one fused forward + backward per shape
            fused_f()
            fused_b()
            torch.cuda.synchronize()
            torch.cuda.cudart().cudaProfilerStart()
            fused_f()
            fused_b()
            torch.cuda.synchronize()
            torch.cuda.cudart().cudaProfilerStop()
            continue
        rec = {"C": C, "HW": HW, "residual": res, "count": cnt, "elems": x.numel()}
        for name, f, b in (("fused", fused_f, fused_b), ("torch", torch_f, fused_b)):
            tf = timeit(f, a.iters, flush)
            f()
            tb = timeit(b, a.iters, flush)
            rec[name + "_fwd_ms"], rec[name + "_bwd_ms"] = round(tf, 4), round(tb, 4)
            tot[name] += cnt * (tf + tb)
        # forward + backward replayed from a CUDA graph: what the layer costs inside the captured training step
        for name, f in (("fused", fused_f), ("torch", torch_f)):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                f()
                fused_b()
            torch.cuda.current_stream().wait_stream(side)
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                f()
                keep = torch.autograd.grad(out["y"], [x, bn.weight, bn.bias] + ([r] if res else []), dy)
            rec[name + "_graph_ms"] = round(timeit(gr.replay, a.iters, flush), 4)
            tot[name + "_graph"] = tot.get(name + "_graph", 0.0) + cnt * rec[name + "_graph_ms"]
            del gr, keep
        if settings.BN_STAMPS:      # GK_BN_STAMPS=1: where inside the kernels the time goes (ns from kernel start)
            fused_f()
            fused_b()
            torch.cuda.synchronize()
            rec["stamps_fwd_bwd_ns"] = phase_stamps(bn)
        n = x.numel() * 4
        fb =n * (3 + (1 if res else 0))                       # stats read + apply read(+res) + write
        bb = n * ((3 if res else 2) + 1 + (3 if res else 3))   # reduce reads (+dres write) + dx reads + write
        rec["fused_fwd_roofline"] = round(fb / bw / (rec["fused_fwd_ms"] * 1e-3), 3)
        rec["fused_bwd_roofline"] = round(bb / bw / (rec["fused_bwd_ms"] * 1e-3), 3)
        rows.append(rec)
        if a.quiet:
            print(C, HW, int(res), "x%d" % cnt, rec["fused_graph_ms"], rec["torch_graph_ms"], flush=True)
        else:
            print(json.dumps(rec), flush=True)
    if a.once:
        return
    summary = {"per_step_ms_fused": round(tot["fused"], 3), "per_step_ms_torch": round(tot["torch"], 3),
               "per_step_graph_ms_fused": round(tot["fused_graph"], 3),
               "per_step_graph_ms_torch": round(tot["torch_graph"], 3),
               "peak_bw_Bps": bw, "batch": a.batch, "rows": rows}
    print(json.dumps({k: v for k, v in summary.items() if k != "rows"}))
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(summary, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    with torch.cuda.stream(torch.cuda.Stream()):   # leaves and graphs on one non-default stream
        main()
