#!/usr/bin/env python
"""ResNet-50 stem convolution, batch 32 x 3 x 224 x 224, forward (and forward + weight gradient): the tcgen05 kernel
(csrc/gk_stem.cu) against cuDNN on the same space-to-depth form and on the original 7x7 / stride 2 form.

    python tools/bench_stem.py [--batch 32] [--iters 30] [--json gpurun_out/stem.json]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200 import settings  # noqa: E402
from gaussiank_sgd_b200.models import fuse  # noqa: E402
from gaussiank_sgd_b200.ops import stem  # noqa: E402


def timed(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.zero_()                       # 256 MB > the 126 MB L2
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    conv = nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
    x = torch.randn(args.batch, 3, args.size, args.size, device=dev).contiguous(memory_format=torch.channels_last)
    res = {"batch": args.batch, "size": args.size}

    def run(fwd_only):
        def f():
            if fwd_only:
                with torch.no_grad():
                    return fuse._stem_conv(conv, x)
            y = fuse._stem_conv(conv, x)
            conv.weight.grad = None
            y.backward(y)                   # dy = y: any dense cotangent
            return y
        return f

    want = None
    for name, tc, s2d, wg in (("cudnn_7x7", False, False, False), ("cudnn_s2d_4x4", False, True, False),
                              ("tcgen05_fwd_cudnn_wgrad", True, True, False), ("tcgen05_s2d_4x4", True, True, True)):
        settings.STEM_TC, settings.STEM_S2D, settings.STEM_TC_WGRAD = tc, s2d, wg
        with torch.no_grad():
            y = fuse._stem_conv(conv, x).float()
        if want is None:
            want = torch.nn.functional.conv2d(x.double(), conv.weight.double(), None, 2, 3)
        res[name] = {"fwd_us": timed(run(True), args.iters), "fwd_wgrad_us": timed(run(False), args.iters),
                     "max_abs_err_vs_fp64": float((y.detach().double() - want.detach()).abs().max())}
    # the implicit-GEMM kernel alone (no space-to-depth kernel, no weight packing)
    settings.STEM_TC = settings.STEM_S2D = settings.STEM_TC_WGRAD = True
    from gaussiank_sgd_b200.ops import native
    lib = native.cuda_lib()
    N, H, W = args.batch, args.size, args.size
    st = torch.cuda.current_stream().cuda_stream
    X = torch.empty(N, H // 2 + 3, W // 2 + 3, 12, device=dev)
    wpk = torch.empty(24 * 512, device=dev)
    y = torch.empty(N, H // 2, W // 2, 64, device=dev)
    e = torch.zeros(1, dtype=torch.int32, device=dev)
    res["s2d_kernel_us"] = timed(lambda: lib.gk_stem_s2d(x.data_ptr(), X.data_ptr(), N, H, W, st), args.iters)
    lib.gk_stem_pack_w(conv.weight.data_ptr(), wpk.data_ptr(), st)
    res["kernel_only_us"] = timed(lambda: lib.gk_stem_fwd(X.data_ptr(), wpk.data_ptr(), y.data_ptr(), N, H // 2, W // 2,
                                                          e.data_ptr(), st), args.iters)
    gw = torch.empty(64, 3, 7, 7, device=dev)
    work = torch.empty(lib.gk_stem_wgrad_workspace(), device=dev)
    res["wgrad_kernels_us"] = timed(lambda: lib.gk_stem_wgrad(y.data_ptr(), X.data_ptr(), gw.data_ptr(), work.data_ptr(), N,
                                                            H // 2, W // 2, e.data_ptr(), st), args.iters)
    res["kernel_bytes"] = X.numel() * 4 + y.numel() * 4
    res["kernel_gbs"] = res["kernel_bytes"] / res["kernel_only_us"] / 1e3
    res["errors"] = stem.error_codes()
    print(json.dumps(res, indent=1))
    if args.json:
        with open(args.json, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
