#!/usr/bin/env python
"""Which cuDNN kernel does the stem's weight gradient get, for a few equivalent formulations?  (batch 32, 224x224)"""
import torch
import torch.nn.functional as F

dev = "cuda:0"
torch.backends.cudnn.benchmark = True
N, OH, OW = 32, 112, 112


def timed(fn, iters=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
    return ts[len(ts) // 2]


def wgrad(gy, X, wshape, stride=1, pad=0):
    w = torch.empty(wshape, device=dev).contiguous(memory_format=torch.channels_last)
    return lambda: torch.ops.aten.convolution_backward(gy, X, w, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1,
                                                       [False, True, False])


gy = torch.randn(N, 64, OH, OW, device=dev).contiguous(memory_format=torch.channels_last)
for C in (12, 16):
    X = torch.randn(N, C, OH + 3, OW + 3, device=dev).contiguous(memory_format=torch.channels_last)
    print(f"s2d 4x4 C={C}: {timed(wgrad(gy, X, (64, C, 4, 4))):.1f} us")
x3 = torch.randn(N, 3, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
print(f"7x7/2 C=3: {timed(wgrad(gy, x3, (64, 3, 7, 7), 2, 3)):.1f} us")
x4 = torch.randn(N, 4, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
print(f"7x7/2 C=4: {timed(wgrad(gy, x4, (64, 4, 7, 7), 2, 3)):.1f} us")
x8 = torch.randn(N, 8, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
print(f"7x7/2 C=8: {timed(wgrad(gy, x8, (64, 8, 7, 7), 2, 3)):.1f} us")
# NCHW
gyn = gy.contiguous()
Xn = torch.randn(N, 12, OH + 3, OW + 3, device=dev)
wn = torch.empty(64, 12, 4, 4, device=dev)
print(f"s2d 4x4 C=12 NCHW: {timed(lambda: torch.ops.aten.convolution_backward(gyn, Xn, wn, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [False, True, False])):.1f} us")
