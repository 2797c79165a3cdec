"""Sparse-allgather transports: where the compaction kernel stores its pairs and where the
receive kernel finds everybody's.

The reference issues two Horovod ``allgather_async`` calls per group (values and int32 indexes,
distributed_optimizer.py:426-427) and blocks on them in ``synchronize`` (:461-463).  Here the
"collective" is part of the kernels:

* :class:`P2PTransport` — every rank owns a double-buffered receive arena in peer-mapped device
  memory (``torch.distributed._symmetric_memory`` or, failing that, CUDA IPC handles of a raw
  allocation).  Rank r's compaction kernel stores its pairs straight into slot r of every peer's
  arena over NVLink and publishes {count, seq}; the receive kernel spins (bounded) on the headers.
  No NCCL call on the sparse path.
* :class:`AllGatherTransport` — one fixed-size ``all_gather_into_tensor`` of (header + pairs); the
  NCCL baseline on GPUs and the gloo path on CPUs (BASELINE.json config 1).
* :class:`LocalTransport` — world of one.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native as N
from ..ops.chain import SlotBuffer, make_dst
from . import comm


class Transport:
    name = "base"
    wait = False

    def dst(self) -> N.GkDst:
        raise NotImplementedError

    def src(self) -> N.GkSrc:
        raise NotImplementedError

    def exchange(self) -> None:
        """Host-issued part of the exchange, enqueued after the send kernel (no-op for P2P)."""

    def send_tensor(self) -> torch.Tensor:
        """The local send slot as an int64 tensor (header words + pairs) for senders that are not kernels."""
        raise NotImplementedError(f"{self.name} transport is written by the compaction kernel only")

    def close(self) -> None:
        pass


class LocalTransport(Transport):
    name = "local"

    def __init__(self, cap: int, device):
        self.buf = SlotBuffer(1, cap, device)
        self._dst = make_dst([self.buf.slot_ptr(0)], 0)
        self._src = self.buf.src(wait=False)

    def dst(self):
        return self._dst

    def src(self):
        return self._src

    def send_tensor(self):
        return self.buf.t


class AllGatherTransport(Transport):
    name = "allgather"

    def __init__(self, cap: int, device, group=None):
        self.P = comm.size()
        self.group = group
        self.send = SlotBuffer(1, cap, device)
        self.recv = SlotBuffer(self.P, cap, device)
        self._dst = make_dst([self.send.slot_ptr(0)], 0)
        self._src = self.recv.src(wait=False)
        self._stage = None
        if device.type == "cuda" and dist.is_initialized() and dist.get_backend() == "gloo":
            self._stage = (torch.empty_like(self.send.t, device="cpu"), torch.empty_like(self.recv.t, device="cpu"))

    def dst(self):
        return self._dst

    def src(self):
        return self._src

    def send_tensor(self):
        return self.send.t

    def exchange(self):
        if self.P == 1:
            self.recv.t.copy_(self.send.t)
        elif self._stage is not None:
            s, r = self._stage
            s.copy_(self.send.t)
            dist.all_gather_into_tensor(r, s, group=self.group)
            self.recv.t.copy_(r)
        else:
            dist.all_gather_into_tensor(self.recv.t, self.send.t, group=self.group)


class P2PTransport(Transport):
    """Peer-mapped, double-buffered receive arenas; the kernels do the communication."""
    name = "p2p"
    wait = True

    def __init__(self, cap: int, device, timeout_s: float = 20.0, backend: str = "auto"):
        assert device.type == "cuda"
        self.P, self.rank = comm.size(), comm.rank()
        self.cap = int(cap)
        self.device = device
        self.timeout_s = timeout_s
        slot_stride = N.GK_SLOT_HDR + self.cap
        words = 2 * self.P * slot_stride
        self._ipc_ptrs: List[int] = []
        self._own_raw = None
        self._symm_t = None
        self.backend = None
        errs = []
        # Every attempt ends with a collective vote: a backend is used only if it came up on EVERY rank, so the
        # ranks can never end up on different transports (or in different collectives) after a partial failure.
        if backend in ("auto", "symm"):
            try:
                self._init_symm(words)
                ok = True
            except Exception as e:  # noqa: BLE001 - any failure falls through to CUDA IPC
                errs.append(f"symmetric memory: {type(e).__name__}: {e}")
                ok = False
            if _all_ranks_ok(ok, device):
                self.backend = "symm"
            else:
                self._symm_t = None
                if backend == "symm":
                    raise RuntimeError("symmetric-memory transport did not come up on every rank: " + "; ".join(errs))
        if self.backend is None:
            try:
                self._init_ipc(words)
                ok = True
            except Exception as e:  # noqa: BLE001
                errs.append(f"cuda ipc: {type(e).__name__}: {e}")
                ok = False
            if _all_ranks_ok(ok, device):
                self.backend = "ipc"
            else:
                self._release_ipc()
                raise RuntimeError("P2P transport did not come up on every rank: " + "; ".join(errs))
        self.errors = errs
        self.slot_stride = slot_stride
        parity_stride = self.P * slot_stride
        my_slots = [base + 8 * self.rank * slot_stride for base in self.peer_ptrs]
        self._dst = make_dst(my_slots, parity_stride)
        self._src = N.GkSrc(self.peer_ptrs[self.rank], slot_stride, parity_stride, self.P, 1,
                            int(timeout_s * 1e9))
        comm.barrier()

    def _init_symm(self, words: int) -> None:
        import torch.distributed._symmetric_memory as symm
        group = dist.group.WORLD
        try:
            symm.enable_symm_mem_for_group(group.group_name)
        except Exception:  # noqa: BLE001 - newer torch enables groups implicitly
            pass
        t = symm.empty(words, dtype=torch.int64, device=self.device)
        t.zero_()
        hdl = symm.rendezvous(t, group=group)
        self.peer_ptrs = [int(p) for p in hdl.buffer_ptrs]
        assert len(self.peer_ptrs) == self.P
        self._symm_t, self._symm_h = t, hdl
        torch.cuda.synchronize(self.device)
        hdl.barrier()

    def _init_ipc(self, words: int) -> None:
        lib = N.cuda_lib()
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        N.check(lib.gk_ipc_alloc(words * 8, C.byref(ptr), handle), "gk_ipc_alloc")
        self._own_raw = ptr.value
        handles: List[Optional[bytes]] = [None] * self.P
        dist.all_gather_object(handles, handle.raw)
        ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(self._own_raw)
            else:
                p = C.c_void_p()
                N.check(lib.gk_ipc_open(h, C.byref(p)), f"gk_ipc_open(rank {r})")
                ptrs.append(p.value)
                self._ipc_ptrs.append(p.value)
        self.peer_ptrs = ptrs

    def dst(self):
        return self._dst

    def src(self):
        return self._src

    def local_words(self) -> torch.Tensor:
        """Copy of this rank's receive arena as an int64 tensor (verification / debugging; synchronises)."""
        words = 2 * self.P * self.slot_stride
        if self._symm_t is not None:
            return self._symm_t[:words].clone()
        out = torch.empty(words, dtype=torch.int64, device=self.device)
        N.check(N.cuda_lib().gk_copy(out.data_ptr(), self._own_raw, 2 * words,
                                     torch.cuda.current_stream(self.device).cuda_stream), "gk_copy")
        torch.cuda.synchronize(self.device)
        return out

    def _release_ipc(self):
        lib = N.cuda_lib()
        for p in self._ipc_ptrs:
            lib.gk_ipc_close(p)
        self._ipc_ptrs = []
        if self._own_raw:
            lib.gk_ipc_free(self._own_raw)
            self._own_raw = None

    def close(self):
        if self.backend == "ipc":
            torch.cuda.synchronize(self.device)
            comm.barrier()
            self._release_ipc()


def _all_ranks_ok(ok: bool, device) -> bool:
    """Collective AND of a per-rank success flag (one tiny all-reduce)."""
    if not comm.is_distributed():
        return ok
    dev = device if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t.item()))


def make_transport(cap: int, device, kind: Optional[str] = None) -> Transport:
    """Pick the transport: ``GK_TRANSPORT`` = auto | p2p | symm | ipc | allgather.

    The P2P transport stores into at most ``GK_MAXP`` = 16 peers from one kernel (one NVSwitch domain of a
    B200 node has 8 GPUs); larger worlds, worlds that span several nodes (no NVLink between them) and CPU runs
    use the all-gather transport, whose receive side handles any number of ranks in chunks of 16."""
    kind = (kind or os.environ.get("GK_TRANSPORT", "auto")).lower()
    P = comm.size()
    if P == 1 and kind in ("auto", "local"):
        return LocalTransport(cap, device)
    if device.type != "cuda" or kind == "allgather":
        return AllGatherTransport(cap, device)
    if kind in ("auto", "p2p", "symm", "ipc"):
        multi_node = comm.local_size() < P
        if P > N.GK_MAXP or multi_node:
            if kind != "auto":
                raise ValueError(f"transport {kind!r} needs all {P} ranks on one NVLink domain of at most "
                                 f"{N.GK_MAXP} GPUs (local size {comm.local_size()}); use 'allgather' or 'auto'")
            return AllGatherTransport(cap, device)
        try:
            return P2PTransport(cap, device, backend={"p2p": "auto"}.get(kind, kind),
                                timeout_s=float(os.environ.get("GK_P2P_TIMEOUT_S", "20")))
        except Exception as e:  # noqa: BLE001 - raised on every rank alike (collective vote inside)
            if kind != "auto":
                raise
            import warnings
            warnings.warn(f"P2P transport unavailable ({type(e).__name__}: {e}); falling back to NCCL all_gather")
            return AllGatherTransport(cap, device)
    raise ValueError(f"unknown transport {kind!r}")
