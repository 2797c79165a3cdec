#!/usr/bin/env python
"""Small-shape exercise of every kernel family, meant to be run under compute-sanitizer:

    compute-sanitizer --tool memcheck  python tools/sanitize.py
    compute-sanitizer --tool racecheck python tools/sanitize.py      # shared-memory hazards
    compute-sanitizer --tool synccheck python tools/sanitize.py

(The GPU pool this framework was developed on does not allow compute-sanitizer; there the script only runs as
a plain smoke test of all paths and the tests' comparisons against PyTorch / the C++ twins stand in for it.)

Shapes are tiny (the tools slow kernels down 10-100x) but chosen to reach every code path: cooperative and
two-kernel BN, barrier-free BN, residual / second gradient / pool / pair variants, slabs that do not fit in
shared memory, the compress chain in gaussian and top-k mode, scatter, fused SGD, LARS, gradient norm.
"""
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200 import settings  # noqa: E402
from gaussiank_sgd_b200.ops import chain as ch  # noqa: E402
from gaussiank_sgd_b200.ops.bn import bn2_act, bn_act, bn_relu_maxpool, coop_errors  # noqa: E402


def bn_cases(dev):
    torch.manual_seed(0)
    for coop in (True, False):
        settings.FUSED_BN_COOP = coop
        for (N, C, H, W) in [(2, 32, 6, 5), (3, 128, 4, 4), (2, 512, 2, 2), (2, 12, 3, 3), (40, 64, 24, 24)]:
            for res, relu, twice in [(False, True, False), (True, True, True), (True, False, False)]:
                bn = nn.BatchNorm2d(C).to(dev)
                x = torch.randn(N, C, H, W, device=dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
                r = torch.randn_like(x).requires_grad_(True) if res else None
                out = bn_act(bn, x, residual=r, relu=relu, twice=twice)
                y = out[0] + 2 * out[1] if twice else out
                y.sum().backward()
                assert coop_errors(bn) == 0
    settings.FUSED_BN_COOP = True
    for (N, C, H, W) in [(2, 64, 4, 4), (2, 2048, 2, 2), (24, 64, 32, 32)]:
        a, b = nn.BatchNorm2d(C).to(dev), nn.BatchNorm2d(C).to(dev)
        xa = torch.randn(N, C, H, W, device=dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        xb = torch.randn_like(xa).requires_grad_(True)
        ya, yb = bn2_act(a, xa, b, xb, twice=True)
        (ya.sum() + 3 * yb.sum()).backward()
        assert coop_errors(a) == 0
    for (N, C, H, W) in [(2, 32, 8, 10), (1, 64, 7, 9)]:
        bn = nn.BatchNorm2d(C).to(dev)
        x = torch.randn(N, C, H, W, device=dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        bn_relu_maxpool(bn, x).sum().backward()
        assert coop_errors(bn) == 0


def chain_cases(dev):
    torch.manual_seed(1)
    n = 70000
    for comp in ("gaussian", "topk", "redsync"):
        g = torch.randn(n, device=dev)
        r = torch.randn(n, device=dev) * 0.1
        c = ch.CompressChain(g, r, [[(0, 30000)], [(30000, n)]], compressor=comp, density=0.01)
        sb = ch.SlotBuffer(1, c.cap, dev)
        for _ in range(2):
            g.normal_()
            c.compress(ch.make_dst([sb.slot_ptr(0)], 0))
        torch.cuda.synchronize()
        assert c.read_ctrl()["err"] == 0


def main():
    dev = torch.device("cuda:0")
    bn_cases(dev)
    chain_cases(dev)
    torch.cuda.synchronize()
    print("sanitize.py: all kernel families ran")


if __name__ == "__main__":
    main()
