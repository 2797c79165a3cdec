"""`gk_grad_norm_static`: the gradient-norm entry point for callers that cannot supply scratch memory.

Who needs it: a host program that reserved all device memory at start-up and uses this library only
through the C ABI (it has no allocator to ask), a watchdog / abort path that must still work when the
device is out of memory, or code that runs next to busy streams and must not call cudaMalloc/cudaFree
(both synchronise the device).  The partial sums and the ticket of the norm kernel then live in a fixed
`static __device__` array inside the library (at most GK_NOALLOC_MAX_BLOCKS = 16384 table entries).

The test goes through the C ABI with raw pointers, exactly like such a caller would, and checks the result
against the ordinary entry point (`gk_grad_norm`, caller-provided scratch) on two calls in a row: the array
is shared between calls and keeps the previous call's contents, so the second call proves that the state
that must start at zero really is re-initialised.  It is collected last (file name) so that a run that
stops at the first failure has already executed every other test, and it passes both alone in a fresh
process (first call = symbol look-up) and after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

from gaussiank_sgd_b200.ops import chain as ch
from gaussiank_sgd_b200.ops import native as N

pytestmark = pytest.mark.gpu

GK_NOALLOC_MAX_BLOCKS = 16384
GK_ERR_NOALLOC_TOO_LARGE = -2


def _table(sizes, chunk):
    """Arena offsets (256-byte aligned tensors) and the device block table for them."""
    extents, total = [], 0
    for s in sizes:
        extents.append((total, s))
        total += (s + 63) // 64 * 64
    blk, _ = ch.build_tables([[e] for e in extents], chunk)
    return extents, total, blk


def test_static_scratch_grad_norm_matches_regular_entry_point_twice():
    lib = N.cuda_lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    sizes = [300, 1_000_003, 17, 65_536, 2_500_000]
    extents, total, blk = _table(sizes, chunk=4096)
    nblk = blk.shape[0]
    assert nblk <= GK_NOALLOC_MAX_BLOCKS
    blk_t = torch.from_numpy(blk.copy()).to(dev)
    # scratch for the ordinary entry point, provided by the caller
    part = torch.zeros(nblk, dtype=torch.float64, device=dev)
    ctrl = torch.zeros(ch.CTRL_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    out_ref = torch.zeros(2, dtype=torch.float32, device=dev)
    out_new = torch.zeros(2, dtype=torch.float32, device=dev)
    for call, (scale, max_norm) in enumerate([(1.0, 5.0), (0.01, 0.25)]):
        g = torch.Generator().manual_seed(100 + call)
        grad = torch.zeros(total, dtype=torch.float32)
        for (o, s) in extents:
            grad[o:o + s] = torch.randn(s, generator=g) * scale
        grad = grad.to(dev)
        N.check(lib.gk_grad_norm(grad.data_ptr(), blk_t.data_ptr(), nblk, part.data_ptr(), C.c_float(max_norm),
                                 out_ref.data_ptr(), ctrl.data_ptr(), stream), "gk_grad_norm")
        # the new entry point: inputs and outputs only, nothing allocated by anybody
        N.check(lib.gk_grad_norm_static(grad.data_ptr(), blk_t.data_ptr(), nblk, C.c_float(max_norm),
                                        out_new.data_ptr(), stream), "gk_grad_norm_static")
        torch.cuda.synchronize(dev)
        ref, new = out_ref.cpu(), out_new.cpu()
        exact = torch.cat([grad[o:o + s] for o, s in extents]).double().norm().item()
        assert abs(ref[0].item() - exact) <= 1e-5 * exact                      # the existing path is right ...
        assert torch.allclose(new, ref, rtol=1e-5, atol=1e-7), (call, new, ref)  # ... and the new one agrees
        assert abs(new[1].item() - min(1.0, max_norm / (exact + 1e-6))) <= 1e-5


def test_static_scratch_rejects_tables_beyond_its_capacity():
    lib = N.cuda_lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    grad = torch.ones(64, device=dev)
    blk_t = torch.zeros(4, dtype=torch.int32, device=dev)
    out = torch.full((2,), -7.0, dtype=torch.float32, device=dev)
    rc = lib.gk_grad_norm_static(grad.data_ptr(), blk_t.data_ptr(), GK_NOALLOC_MAX_BLOCKS + 1, C.c_float(1.0),
                                 out.data_ptr(), stream)
    torch.cuda.synchronize(dev)
    assert rc == GK_ERR_NOALLOC_TOO_LARGE
    assert out.cpu().tolist() == [-7.0, -7.0]      # nothing was launched


@pytest.mark.multigpu
def test_static_scratch_on_a_second_device():
    """The array exists once per GPU and its address is looked up per device, with that device current: a call
    on device 1 (after device 0 has cached its own address) must use device 1's copy."""
    lib = N.cuda_lib()
    results = []
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        torch.cuda.set_device(dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        sizes = [70_001, 4_099]
        extents, total, blk = _table(sizes, chunk=4096)
        blk_t = torch.from_numpy(blk.copy()).to(dev)
        grad = torch.zeros(total, dtype=torch.float32)
        g = torch.Generator().manual_seed(7 + d)
        for (o, s) in extents:
            grad[o:o + s] = torch.randn(s, generator=g)
        grad = grad.to(dev)
        out = torch.zeros(2, dtype=torch.float32, device=dev)
        N.check(lib.gk_grad_norm_static(grad.data_ptr(), blk_t.data_ptr(), blk.shape[0], C.c_float(1.0),
                                        out.data_ptr(), stream), "gk_grad_norm_static")
        torch.cuda.synchronize(dev)
        exact = torch.cat([grad[o:o + s] for o, s in extents]).double().norm().item()
        results.append((out[0].item(), exact))
    torch.cuda.set_device(0)
    for got, exact in results:
        assert abs(got - exact) <= 1e-5 * exact
