#!/usr/bin/env python
"""Launch every kernel of csrc/gk_stem.cu a few times on the ResNet-50 stem shape (batch 32, 224 x 224), for

    ncu --set full --clock-control none --import-source on -k regex:k_stem -o /tmp/stem python tools/ncu_stem.py

Run it once without ncu first; it prints what it launched.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200.ops import native  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    lib = native.cuda_lib()
    st = torch.cuda.current_stream().cuda_stream
    N, H, W = 32, 224, 224
    torch.manual_seed(0)
    x = torch.randn(N, 3, H, W, device=dev)
    xl = x.contiguous(memory_format=torch.channels_last)
    w = torch.randn(64, 3, 7, 7, device=dev) * 0.05
    X = torch.empty(N, H // 2 + 3, W // 2 + 3, 12, device=dev)
    wpk = torch.empty(24 * 512, device=dev)
    y = torch.empty(N, H // 2, W // 2, 64, device=dev)
    gy = torch.randn_like(y)
    sums = torch.empty(lib.gk_stem_stats_rows(), 128, device=dev)
    work = torch.empty(lib.gk_stem_wgrad_workspace(), device=dev)
    gw = torch.empty(64, 3, 7, 7, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        flush.zero_()
        native.check(lib.gk_stem_s2d_layout(xl.data_ptr(), X.data_ptr(), N, H, W, 0, st))
        native.check(lib.gk_stem_s2d_layout(x.data_ptr(), X.data_ptr(), N, H, W, 1, st))
        native.check(lib.gk_stem_pack_w(w.data_ptr(), wpk.data_ptr(), st))
        flush.zero_()
        native.check(lib.gk_stem_fwd(X.data_ptr(), wpk.data_ptr(), y.data_ptr(), N, H // 2, W // 2, err.data_ptr(), st))
        flush.zero_()
        native.check(lib.gk_stem_fwd_stats(X.data_ptr(), wpk.data_ptr(), y.data_ptr(), sums.data_ptr(), N, H // 2, W // 2,
                                           err.data_ptr(), st))
        flush.zero_()
        native.check(lib.gk_stem_wgrad(gy.data_ptr(), X.data_ptr(), gw.data_ptr(), work.data_ptr(), N, H // 2, W // 2,
                                       err.data_ptr(), st))
    torch.cuda.synchronize()
    print("launched s2d (nhwc, nchw), pack_w, fwd, fwd+stats, wgrad + fold; error word", int(err))


if __name__ == "__main__":
    main()
