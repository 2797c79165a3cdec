"""Python driver of the native kernel chain: block/segment tables, scratch, launches.

A :class:`CompressChain` owns everything one sparsification group needs on one device and
launches the send-side chain (K1 accumulate+moments, K2 count+pick or K4 radix select,
K3 compact+send) without any host synchronisation.  The same object drives the C++ CPU twin
when its tensors live on the CPU (gloo mode).

Reference behaviour implemented: compression.py:358-389 (gaussian), :405-435 (gaussian2),
:45-49 (exact top-k), :52-56 (min-k), :644-677 (redsync), :707-738 (redsynctrim),
distributed_optimizer.py:416-430 (send side).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import native as N

SEG_DTYPE = np.dtype([("n", "<i4"), ("k", "<i4"), ("blk_begin", "<i4"), ("blk_end", "<i4"),
                      ("zabs", "<f8"), ("param", "<f4"), ("aux", "<i4")])
STAT_DTYPE = np.dtype([("mean", "<f4"), ("std", "<f4"), ("meanabs", "<f4"), ("maxabs", "<f4"),
                       ("cand", "<f4", (N.GK_NC,)), ("thr", "<f4"), ("tbits", "<u4"), ("ties", "<u4"),
                       ("chosen", "<i4"), ("count", "<u4"), ("out_base", "<u4"), ("prefix", "<u4"),
                       ("krem", "<u4"), ("rcand", "u1", (N.GK_NC,)), ("nrc", "<i4"),
                       ("tot", "<u4", (N.GK_NC,)), ("known", "<u4"), ("gate", "<i4"), ("done", "<i4"),
                       ("sprefix", "<u4"), ("skrem", "<u4")])
CTRL_DTYPE = np.dtype([("done", "<u4", (8,)), ("total", "<u4"), ("sent", "<u4"), ("seq", "<u4"),
                       ("err", "<u4"), ("overflow", "<u4"), ("rsv", "<u4", (3,)), ("sent_total", "<u8"),
                       ("ring", "<u4", (64,))])
HYPER_DTYPE = np.dtype([("lr", "<f4"), ("momentum", "<f4"), ("dampening", "<f4"), ("weight_decay", "<f4"),
                        ("nesterov", "<i4"), ("first_step", "<i4"), ("lars_eeta", "<f4"), ("lars_eps", "<f4")])
assert SEG_DTYPE.itemsize == 32 and STAT_DTYPE.itemsize == 216 and CTRL_DTYPE.itemsize == 328
assert HYPER_DTYPE.itemsize == 32

DEFAULT_CHUNK = 8192

# kernels of THIS library launched so far (Python-side count; a replayed CUDA graph re-launches the
# kernels counted while it was captured)
LAUNCHES = {"n": 0}

MODES = {
    "gaussian": (N.MODE_GAUSSIAN, 3), "gaussiank": (N.MODE_GAUSSIAN, 3), "gaussian2": (N.MODE_GAUSSIAN, 5),
    "topk": (N.MODE_TOPK, 0), "topk2": (N.MODE_TOPK, 0), "mink": (N.MODE_MINK, 0),
    "redsync": (N.MODE_REDSYNC, 0), "redsynctrim": (N.MODE_REDSYNCTRIM, 0),
    "threshold": (N.MODE_THRESHOLD, 0), "dgcsampling": (N.MODE_DGC, 0),
}
ROW = 128   # elements per row of the row-maximum table (csrc/gk_cuda.cu: GK_ROW)


def compute_k(n: int, density: float) -> int:
    """k = max(int(numel*ratio), 1) in Python double arithmetic (compression.py:363)."""
    return max(int(n * density), 1)


def gaussian_zabs(density: float) -> float:
    """|z| with right_thres = mu + |z| sigma; same expression as utils.py:136-138 called with
    p_value = 1 - ratio (compression.py:369)."""
    from scipy import stats
    p_value = 1 - density
    return float(-stats.norm.ppf((1 - p_value) / 2))


def build_tables(segments: Sequence[Sequence[Tuple[int, int]]], chunk: int = DEFAULT_CHUNK):
    """segments[s] = [(arena_start, numel), ...] extents (tensors) of selection segment s.
    Returns (blk int32[nblk,4], seg structured[nseg]).  Every block is a contiguous extent of at most
    ``chunk`` elements starting on a multiple of 4 (tensors start on aligned arena offsets)."""
    assert chunk % ROW == 0 and chunk > 0
    blks: List[Tuple[int, int, int, int]] = []
    segs = np.zeros(len(segments), dtype=SEG_DTYPE)
    for s, extents in enumerate(segments):
        begin = len(blks)
        soff = 0
        for (start, numel) in extents:
            assert start % 4 == 0, "tensor extents must start 16-byte aligned"
            off = 0
            while off < numel:
                ln = min(chunk, numel - off)
                blks.append((start + off, ln, s, soff))
                off += ln
                soff += ln
        segs[s]["n"] = soff
        segs[s]["blk_begin"] = begin
        segs[s]["blk_end"] = len(blks)
    blk = np.asarray(blks, dtype=np.int32).reshape(-1, 4)
    return blk, segs


def _to_dev(arr: np.ndarray, device) -> torch.Tensor:
    t = torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy())
    return t.to(device)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream_ptr(device) -> Optional[int]:
    if device.type != "cuda":
        return None
    return torch.cuda.current_stream(device).cuda_stream


class SlotBuffer:
    """A receive arena: ``parities`` x ``nslots`` slots of (GK_SLOT_HDR + cap) u64 words."""

    def __init__(self, nslots: int, cap: int, device, parities: int = 1, tensor: Optional[torch.Tensor] = None):
        self.nslots, self.cap, self.parities = nslots, int(cap), parities
        self.slot_stride = N.GK_SLOT_HDR + self.cap
        words = parities * nslots * self.slot_stride
        if tensor is None:
            tensor = torch.zeros(words, dtype=torch.int64, device=device)
        assert tensor.numel() >= words and tensor.dtype == torch.int64
        self.t = tensor
        self.device = tensor.device

    @property
    def parity_stride(self) -> int:
        return self.nslots * self.slot_stride if self.parities > 1 else 0

    def slot_ptr(self, r: int, base_ptr: Optional[int] = None) -> int:
        base = self.t.data_ptr() if base_ptr is None else base_ptr
        return base + 8 * r * self.slot_stride

    def src(self, wait: bool = False, timeout_s: float = 20.0) -> N.GkSrc:
        return N.GkSrc(self.t.data_ptr(), self.slot_stride, self.parity_stride, self.nslots,
                       1 if wait else 0, int(timeout_s * 1e9))

    def read_slot(self, r: int, parity: int = 0):
        """(indices int64, values fp32) of slot r — host-side decode for tests and the compat API."""
        base = (parity * self.nslots + r) * self.slot_stride if self.parities > 1 else r * self.slot_stride
        hdr = int(self.t[base].item())
        cnt = min(hdr & 0xFFFFFFFF, self.cap)
        words = self.t[base + N.GK_SLOT_HDR: base + N.GK_SLOT_HDR + cnt].cpu().numpy().view(np.uint64)
        idx = (words & np.uint64(0xFFFFFFFF)).astype(np.int64)
        val = (words >> np.uint64(32)).astype(np.uint32).view(np.float32)
        return torch.from_numpy(idx), torch.from_numpy(val.copy()), hdr >> 32


def make_dst(slot_ptrs: Sequence[int], parity_stride: int) -> N.GkDst:
    d = N.GkDst()
    assert 1 <= len(slot_ptrs) <= N.GK_MAXP
    for i, p in enumerate(slot_ptrs):
        d.slot[i] = p
    d.ndst = len(slot_ptrs)
    d.parity_stride = parity_stride
    return d


class CompressChain:
    """Send-side chain of one group of selection segments on one device."""

    def __init__(self, grad: torch.Tensor, resid: torch.Tensor,
                 segments: Sequence[Sequence[Tuple[int, int]]], *, compressor: str = "gaussian",
                 density: float = 0.001, use_residual: bool = True, zero_grad: bool = True,
                 cap: Optional[int] = None, chunk: int = DEFAULT_CHUNK, max_density: Optional[float] = None):
        assert grad.dtype == torch.float32 and resid.dtype == torch.float32
        assert grad.is_contiguous() and resid.is_contiguous() and grad.numel() == resid.numel()
        assert grad.numel() < 2 ** 31, "int32 indices on the wire (distributed_optimizer.py:427)"
        self.device = grad.device
        self.is_cuda = self.device.type == "cuda"
        self.grad, self.resid = grad, resid
        self.compressor = compressor
        self.mode, self.nloops = MODES[compressor]
        self.use_residual, self.zero_grad = use_residual, zero_grad
        self._segments = [list(e) for e in segments]
        self.chunk = chunk
        blk, segs = build_tables(segments, chunk)
        self.blk_np, self.seg_np = blk, segs
        self.nblk, self.nseg = blk.shape[0], segs.shape[0]
        self.n_total = int(segs["n"].sum())
        dev = self.device
        # Every table and scratch array of the chain lives in ONE allocation (views carved on 256-byte
        # boundaries): the kernels of a step run with cold TLBs, and each separate allocation would cost the
        # first warp that touches it a page walk in the middle of a latency-bound kernel.
        need_hist = self.mode in (N.MODE_TOPK, N.MODE_MINK, N.MODE_DGC)
        self.rpb = (chunk + ROW - 1) // ROW   # rows of the row-maximum table reserved per table entry
        nb = max(self.nblk, 1)
        layout = [("blk", nb * 16, torch.int32), ("seg", self.nseg * SEG_DTYPE.itemsize, torch.uint8),
                  ("part", nb * 4 * 8, torch.float64), ("stat", self.nseg * STAT_DTYPE.itemsize, torch.uint8),
                  ("bcnt", nb * N.GK_NC * 4, torch.int32), ("boff", nb * 2 * 4, torch.int32),
                  ("rowmax", nb * self.rpb * 4, torch.int32),
                  ("hist", (self.nseg * N.GK_HIST_BINS if need_hist else 1) * 4, torch.int32),
                  ("ctrl", CTRL_DTYPE.itemsize, torch.uint8)]
        off, offs = 0, {}
        for name, nbytes, _dt in layout:
            offs[name] = off
            off += (nbytes + 255) // 256 * 256
        self._scratch = torch.zeros(off + 256, dtype=torch.uint8, device=dev)
        shift = (-self._scratch.data_ptr()) % 256
        for name, nbytes, dt in layout:
            view = self._scratch[shift + offs[name]: shift + offs[name] + nbytes]
            setattr(self, name, view if dt == torch.uint8 else view.view(dt))
        # largest |acc| bit pattern of every 128-element row of every table entry: self.rowmax (written by K1,
        # read by the later passes to skip rows that cannot hold a selected element)
        self.blk.copy_(torch.from_numpy(np.ascontiguousarray(blk).reshape(-1)))
        self._seg_pinned = None
        self.density = None
        dmax = max(density, max_density or density)
        self.auto_cap = cap is None      # capacity follows the density (cap_for); an explicit cap is kept
        if cap is None:
            cap = self.cap_for(dmax)
        self.cap = int(cap)
        self.set_density(density)
        self._lib = N.cuda_lib() if self.is_cuda else N.cpu_lib()
        self._desc = self._make_desc()

    # ------------------------------------------------------------------ configuration
    def cap_for(self, density: float) -> int:
        """Receive-slot capacity for a density: 8x k (the gaussian estimate selects up to
        ~10x k on skewed residuals); pairs beyond it stay in the residual (error feedback keeps them)."""
        k = sum(compute_k(int(n), density) for n in self.seg_np["n"])
        return int(min(max(8 * k, 1024), max(self.n_total, 1024)))

    def set_cap(self, cap: int) -> None:
        """New slot capacity (the caller re-creates the transport to match)."""
        self.cap = int(cap)
        self._desc.cap = self.cap

    def set_density(self, density: float, threshold_param: Optional[float] = None) -> None:
        """Rewrite k / zabs of every segment (only when the density changes, i.e. per epoch:
        distributed_optimizer.py:146-160)."""
        if density == self.density and threshold_param is None:
            return
        self.density = float(density)
        segs = self.seg_np
        kb = 0
        z = gaussian_zabs(density) if self.mode == N.MODE_GAUSSIAN and density < 1 else 0.0
        for s in range(self.nseg):
            k = compute_k(int(segs[s]["n"]), density)
            segs[s]["k"] = k
            segs[s]["zabs"] = z
            segs[s]["aux"] = kb
            if threshold_param is not None:
                segs[s]["param"] = threshold_param
            kb += k
        self.k_total = kb
        host = torch.from_numpy(segs.view(np.uint8).reshape(-1).copy())
        if self.is_cuda:
            if self._seg_pinned is None:
                self._seg_pinned = torch.empty_like(host).pin_memory()
            self._seg_pinned.copy_(host)
            self.seg.copy_(self._seg_pinned, non_blocking=True)
        else:
            self.seg.copy_(host)

    # ------------------------------------------------------------- external gradient sources
    def _extent_of_blocks(self):
        """For every table entry: (index of its tensor extent in flattened segment order, element offset
        inside that extent)."""
        if getattr(self, "_blk_extent", None) is None:
            ext, off, e = [], [], 0
            for extents in self._segments:
                for (_start, numel) in extents:
                    o = 0
                    while o < numel:
                        ext.append(e)
                        off.append(o)
                        o += min(self.chunk, numel - o)
                    e += 1
            self._blk_extent = np.asarray(ext, dtype=np.int64)
            self._blk_eoff = np.asarray(off, dtype=np.int64)
            self.n_extents = e
        return self._blk_extent, self._blk_eoff

    def set_grad_sources(self, extent_ptrs: Optional[Sequence[int]]) -> None:
        """Make K1 read the gradient of tensor extent e from device address ``extent_ptrs[e]`` (flattened
        segment order) instead of the gradient arena — e.g. the tensors autograd just produced.  The table
        of per-entry pointers is staged in rotating pinned buffers and copied on the current stream, so it
        may be refreshed every iteration (eager mode) or captured once (CUDA graph).  ``None`` switches back
        to the arena."""
        if extent_ptrs is None:
            self._desc.gsrc = None
            return
        be, bo = self._extent_of_blocks()
        ptrs = np.asarray(extent_ptrs, dtype=np.int64)
        assert ptrs.shape[0] == self.n_extents
        table = ptrs[be] + 4 * bo
        if not self.is_cuda:
            self._gsrc_dev = torch.from_numpy(table.copy())
            self._desc.gsrc = self._gsrc_dev.data_ptr()
            return
        if getattr(self, "_gsrc_dev", None) is None:
            self._gsrc_dev = torch.zeros(max(self.nblk, 1), dtype=torch.int64, device=self.device)
            self._gsrc_pin = [torch.zeros(max(self.nblk, 1), dtype=torch.int64).pin_memory() for _ in range(4)]
            self._gsrc_evt = [None] * 4
            self._gsrc_i = 0
        i = self._gsrc_i
        self._gsrc_i = (i + 1) % 4
        if self._gsrc_evt[i] is not None:
            self._gsrc_evt[i].synchronize()      # the copy that last used this staging buffer has run
        self._gsrc_pin[i].numpy()[:table.shape[0]] = table
        self._gsrc_dev.copy_(self._gsrc_pin[i], non_blocking=True)
        if not torch.cuda.is_current_stream_capturing():
            ev = torch.cuda.Event()
            ev.record()
            self._gsrc_evt[i] = ev
        self._desc.gsrc = self._gsrc_dev.data_ptr()

    def _make_desc(self) -> N.GkChain:
        return N.GkChain(self.grad.data_ptr(), self.resid.data_ptr(), self.blk.data_ptr(), self.seg.data_ptr(),
                         self.part.data_ptr(), self.stat.data_ptr(), self.bcnt.data_ptr(), self.boff.data_ptr(),
                         self.hist.data_ptr(), self.ctrl.data_ptr(), self.nblk, self.nseg, self.mode,
                         self.nloops, 1 if self.use_residual else 0, 1 if self.zero_grad else 0, self.cap, 0,
                         None, self.rowmax.data_ptr() if self.is_cuda else None, self.rpb, 0)

    # ------------------------------------------------------------------------ launches
    def compress(self, dst: N.GkDst, seed: int = 0) -> None:
        """Enqueue the whole send-side chain on the current stream (no host sync).  ``seed`` keys the
        coordinate sample of the DGC mode (compression.py:588-591 draws a fresh permutation every call)."""
        if self.mode == N.MODE_DGC:
            LAUNCHES["n"] += 9
            if self.is_cuda:
                N.check(self._lib.gk_dgc_compress(C.byref(self._desc), C.byref(dst), C.c_uint64(seed),
                                                  _stream_ptr(self.device)), "gk_dgc_compress")
            else:
                N.check(self._lib.gkcpu_dgc_compress(C.byref(self._desc), C.byref(dst), C.c_uint64(seed)),
                        "gkcpu_dgc_compress")
            return
        LAUNCHES["n"] += 5 if self.mode in (N.MODE_TOPK, N.MODE_MINK) else 3
        if self.is_cuda:
            N.check(self._lib.gk_compress(C.byref(self._desc), C.byref(dst), _stream_ptr(self.device)),
                    "gk_compress")
        else:
            N.check(self._lib.gkcpu_compress(C.byref(self._desc), C.byref(dst)), "gkcpu_compress")

    def accumulate(self) -> None:
        assert self.is_cuda
        N.check(self._lib.gk_accumulate(C.byref(self._desc), _stream_ptr(self.device)), "gk_accumulate")

    def count_pick(self) -> None:
        assert self.is_cuda
        N.check(self._lib.gk_count_pick(C.byref(self._desc), _stream_ptr(self.device)), "gk_count_pick")

    def topk_select(self) -> None:
        assert self.is_cuda
        N.check(self._lib.gk_topk_select(C.byref(self._desc), _stream_ptr(self.device)), "gk_topk_select")

    def compact_send(self, dst: N.GkDst) -> None:
        assert self.is_cuda
        N.check(self._lib.gk_compact_send(C.byref(self._desc), C.byref(dst), _stream_ptr(self.device)),
                "gk_compact_send")

    def randomk_send(self, dst: N.GkDst, seed: int) -> None:
        """acc -> resid, then k distinct pseudo-random coordinates per segment
        (compression.py:461-475 / :497-509)."""
        LAUNCHES["n"] += 2
        if self.is_cuda:
            st = _stream_ptr(self.device)
            N.check(self._lib.gk_accumulate(C.byref(self._desc), st), "gk_accumulate")
            N.check(self._lib.gk_randomk_send(C.byref(self._desc), C.byref(dst), C.c_uint64(seed), st),
                    "gk_randomk_send")
        else:
            self._cpu_accumulate_only()
            N.check(self._lib.gkcpu_randomk_send(C.byref(self._desc), C.byref(dst), C.c_uint64(seed)),
                    "gkcpu_randomk_send")

    def _cpu_accumulate_only(self) -> None:
        for (start, ln, _s, _o) in self.blk_np:
            g = self.grad[start:start + ln]
            r = self.resid[start:start + ln]
            if self.use_residual:
                r.add_(g)
            else:
                r.copy_(g)
            if self.zero_grad:
                g.zero_()

    def bucket_means(self, means: torch.Tensor) -> None:
        LAUNCHES["n"] += 1
        if self.is_cuda:
            N.check(self._lib.gk_bucket_means(C.byref(self._desc), means.data_ptr(), _stream_ptr(self.device)))
        else:
            N.check(self._lib.gkcpu_bucket_means(C.byref(self._desc), means.data_ptr()))

    def bucket_reconstruct(self, local_means: torch.Tensor, global_means: torch.Tensor) -> None:
        LAUNCHES["n"] += 1
        if self.is_cuda:
            N.check(self._lib.gk_bucket_reconstruct(C.byref(self._desc), local_means.data_ptr(),
                                                    global_means.data_ptr(), _stream_ptr(self.device)))
        else:
            N.check(self._lib.gkcpu_bucket_reconstruct(C.byref(self._desc), local_means.data_ptr(),
                                                       global_means.data_ptr()))

    # --------------------------------------------------------------- host-side readback
    def read_ctrl(self) -> dict:
        a = self.ctrl.cpu().numpy().view(CTRL_DTYPE)[0]
        return {k: (a[k].tolist() if a[k].shape else a[k].item()) for k in CTRL_DTYPE.names}

    def read_stats(self) -> np.ndarray:
        return self.stat.cpu().numpy().view(STAT_DTYPE).copy()


# ------------------------------------------------------------------- receive side / optimizer
def _src_chunks(src: N.GkSrc):
    """The receive kernels take at most GK_MAXP slots per launch: split a wider arena into rank-ordered views."""
    if src.nsrc <= N.GK_MAXP:
        yield src
        return
    for r0 in range(0, src.nsrc, N.GK_MAXP):
        yield N.GkSrc(src.arena + 8 * r0 * src.slot_stride, src.slot_stride, src.parity_stride,
                      min(N.GK_MAXP, src.nsrc - r0), src.wait, src.timeout_ns)


def scatter(src: N.GkSrc, ctrl: torch.Tensor, grad: torch.Tensor, scale: float) -> None:
    """grad[idx] += val*scale over every slot, in rank order (distributed_optimizer.py:468-482): one launch
    for up to two sources, one launch per source rank beyond that (bit-identical sums on every replica)."""
    LAUNCHES["n"] += src.nsrc if (src.nsrc > 2 and grad.device.type == "cuda") else 1
    for part in _src_chunks(src):
        if grad.device.type == "cuda":
            N.check(N.cuda_lib().gk_scatter(C.byref(part), ctrl.data_ptr(), grad.data_ptr(), C.c_float(scale),
                                            grad.numel(), _stream_ptr(grad.device)), "gk_scatter")
        else:
            N.check(N.cpu_lib().gkcpu_scatter(C.byref(part), ctrl.data_ptr(), grad.data_ptr(), C.c_float(scale),
                                              grad.numel()), "gkcpu_scatter")


def unscatter(src: N.GkSrc, ctrl: torch.Tensor, grad: torch.Tensor) -> None:
    LAUNCHES["n"] += 1
    for part in _src_chunks(src):
        if grad.device.type == "cuda":
            N.check(N.cuda_lib().gk_unscatter(C.byref(part), ctrl.data_ptr(), grad.data_ptr(), grad.numel(),
                                              _stream_ptr(grad.device)), "gk_unscatter")
        else:
            N.check(N.cpu_lib().gkcpu_unscatter(C.byref(part), ctrl.data_ptr(), grad.data_ptr(), grad.numel()),
                    "gkcpu_unscatter")


class FlatTables:
    """Block table over a flat arena for the optimizer-side kernels; ``seg`` of every block is the
    index returned by ``group_of(tensor_index)``."""

    def __init__(self, extents: Sequence[Tuple[int, int]], group_of: Sequence[int], device,
                 chunk: int = DEFAULT_CHUNK):
        blk, _ = build_tables([[e] for e in extents], chunk)
        self.tensor_blk = blk.copy()                      # seg = tensor index
        gblk = blk.copy()
        gblk[:, 2] = np.asarray(group_of, dtype=np.int32)[blk[:, 2]]
        self.nblk = blk.shape[0]
        self.group_blk_t = _to_dev(gblk, device)          # seg = param group (SGD)
        self.tensor_blk_t = _to_dev(blk, device)          # seg = tensor (LARS)
        tens = np.zeros(len(extents), dtype=SEG_DTYPE)
        b = 0
        for t, (_s, numel) in enumerate(extents):
            nb = (numel + chunk - 1) // chunk
            tens[t]["n"] = numel
            tens[t]["blk_begin"] = b
            tens[t]["blk_end"] = b + nb
            tens[t]["aux"] = group_of[t]
            b += nb
        self.ntens = len(extents)
        self.tens_t = _to_dev(tens, device)
        self.part = torch.zeros(max(self.nblk, 1) * 2, dtype=torch.float64, device=device)
        self.ratio = torch.zeros(max(self.ntens, 1), dtype=torch.float32, device=device)
        self.norm_out = torch.ones(2, dtype=torch.float32, device=device)
        self.ctrl = torch.zeros(CTRL_DTYPE.itemsize, dtype=torch.uint8, device=device)
        self.device = torch.device(device)
        # folded receive+update (gk_sgd_recv): param group of every 64-element granule of the arena (tensors
        # start on 64-element boundaries, parallel/arena.py) and the grid-barrier words
        total = max((s + n + 63) // 64 for s, n in extents) if len(extents) else 1
        gmap = np.zeros(total, dtype=np.uint8)
        self.foldable = max(group_of, default=0) < 256 and all(s % 64 == 0 for s, _ in extents)
        if self.foldable:
            for (s, n), g in zip(extents, group_of):
                gmap[s // 64:(s + n + 63) // 64] = g
        self.gmap = torch.from_numpy(gmap).to(device)
        self.bar = torch.zeros(4, dtype=torch.int32, device=device)


def sgd_step(param: torch.Tensor, grad: torch.Tensor, mom: Optional[torch.Tensor], tables: FlatTables,
             hyper: torch.Tensor, grad_scale: Optional[torch.Tensor] = None) -> None:
    """One fused momentum-SGD pass over the arenas (torch.optim.SGD math, dl_trainer.py:848-872)."""
    LAUNCHES["n"] += 1
    if param.device.type == "cuda":
        N.check(N.cuda_lib().gk_sgd(param.data_ptr(), grad.data_ptr(), _ptr(mom), tables.group_blk_t.data_ptr(),
                                    tables.nblk, hyper.data_ptr(), _ptr(grad_scale),
                                    _stream_ptr(param.device)), "gk_sgd")
    else:
        N.check(N.cpu_lib().gkcpu_sgd(param.data_ptr(), grad.data_ptr(), _ptr(mom), tables.group_blk_t.data_ptr(),
                                      tables.nblk, hyper.data_ptr(), _ptr(grad_scale)), "gkcpu_sgd")


def sgd_recv_step(param: torch.Tensor, mask: torch.Tensor, mom: Optional[torch.Tensor], tables: FlatTables,
                  hyper: torch.Tensor, src: N.GkSrc, ctrl: torch.Tensor, scale: float,
                  skip_dense: bool = False) -> None:
    """Folded receive + momentum-SGD (include/gk_b200.h: gk_sgd_recv): the dense weight-decay / momentum pass
    runs while the peers' pairs are in flight, then the received pairs are folded in per coordinate in rank
    order; no dense gradient is formed.  ``mask`` is the all-zero gradient arena (scratch, zero again
    afterwards).  Replaces distributed_optimizer.py:461-482 + :543-546 for the common path."""
    assert tables.foldable
    LAUNCHES["n"] += 1
    u = N.GkRecvUpd(param.data_ptr(), _ptr(mom), mask.data_ptr(), tables.group_blk_t.data_ptr(),
                    hyper.data_ptr(), tables.gmap.data_ptr(), ctrl.data_ptr(), tables.bar.data_ptr(),
                    float(scale), tables.nblk, mask.numel(), 1 if skip_dense else 0)
    if param.device.type == "cuda":
        N.check(N.cuda_lib().gk_sgd_recv(C.byref(u), C.byref(src), _stream_ptr(param.device)), "gk_sgd_recv")
    else:
        N.check(N.cpu_lib().gkcpu_sgd_recv(C.byref(u), C.byref(src)), "gkcpu_sgd_recv")


def lars_step(param: torch.Tensor, grad: torch.Tensor, acc: torch.Tensor, tables: FlatTables,
              hyper: torch.Tensor) -> None:
    """LARS (lars.py:67-134): segmented norms, trust ratio, clamped update."""
    LAUNCHES["n"] += 2
    if param.device.type == "cuda":
        N.check(N.cuda_lib().gk_lars(param.data_ptr(), grad.data_ptr(), acc.data_ptr(),
                                     tables.tensor_blk_t.data_ptr(), tables.nblk, tables.tens_t.data_ptr(),
                                     tables.ntens, hyper.data_ptr(), tables.part.data_ptr(),
                                     tables.ratio.data_ptr(), tables.ctrl.data_ptr(),
                                     _stream_ptr(param.device)), "gk_lars")
    else:
        N.check(N.cpu_lib().gkcpu_lars(param.data_ptr(), grad.data_ptr(), acc.data_ptr(),
                                       tables.tensor_blk_t.data_ptr(), tables.nblk, tables.tens_t.data_ptr(),
                                       tables.ntens, hyper.data_ptr(), tables.part.data_ptr(),
                                       tables.ratio.data_ptr()), "gkcpu_lars")


def grad_norm(grad: torch.Tensor, tables: FlatTables, max_norm: float) -> torch.Tensor:
    """Global L2 norm and clip coefficient on the device: out = [norm, coef] (dist_trainer.py:80-85)."""
    LAUNCHES["n"] += 1
    if grad.device.type == "cuda":
        N.check(N.cuda_lib().gk_grad_norm(grad.data_ptr(), tables.group_blk_t.data_ptr(), tables.nblk,
                                          tables.part.data_ptr(), C.c_float(max_norm),
                                          tables.norm_out.data_ptr(), tables.ctrl.data_ptr(),
                                          _stream_ptr(grad.device)), "gk_grad_norm")
    else:
        N.check(N.cpu_lib().gkcpu_grad_norm(grad.data_ptr(), tables.group_blk_t.data_ptr(), tables.nblk,
                                            C.c_float(max_norm), tables.norm_out.data_ptr()), "gkcpu_grad_norm")
    return tables.norm_out


def scale_(grad: torch.Tensor, tables: FlatTables, scale: torch.Tensor) -> None:
    LAUNCHES["n"] += 1
    if grad.device.type == "cuda":
        N.check(N.cuda_lib().gk_scale(grad.data_ptr(), tables.group_blk_t.data_ptr(), tables.nblk,
                                      scale.data_ptr(), _stream_ptr(grad.device)), "gk_scale")
    else:
        s = float(scale.reshape(-1)[0])
        if s != 1.0:
            grad.mul_(s)


def make_hyper(groups: Sequence[dict], device, first_step: bool) -> torch.Tensor:
    arr = np.zeros(len(groups), dtype=HYPER_DTYPE)
    for i, g in enumerate(groups):
        arr[i]["lr"] = g["lr"]
        arr[i]["momentum"] = g.get("momentum", 0.0)
        arr[i]["dampening"] = g.get("dampening", 0.0)
        arr[i]["weight_decay"] = g.get("weight_decay", 0.0)
        arr[i]["nesterov"] = 1 if g.get("nesterov", False) else 0
        arr[i]["first_step"] = 1 if first_step else 0
        arr[i]["lars_eeta"] = g.get("eeta", 0.001)
        arr[i]["lars_eps"] = g.get("epsilon", 1e-5)
    return torch.from_numpy(arr.view(np.uint8).reshape(-1).copy())
