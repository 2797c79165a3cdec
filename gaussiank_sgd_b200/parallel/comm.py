"""Process-group plumbing with the Horovod-style surface the reference programs against.

The reference imports ``init/size/rank/local_rank/local_size/broadcast/synchronize`` and the async
collectives from ``horovod.torch.mpi_ops`` (distributed_optimizer.py:21-26) and launches with
``mpirun`` (run.sh:13-15).  Here there is one process per GPU started by ``torchrun`` /
``torch.multiprocessing.spawn``; ``torch.distributed`` (NCCL on GPUs, gloo on CPUs) is the control
plane and the dense baseline, and the sparse path goes through ``parallel.transport``.
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist

_state = {"initialized": False, "owns_pg": False, "device": None}


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def init(backend: Optional[str] = None, device: Optional[torch.device] = None, timeout_s: int = 600) -> None:
    """Join the job.  World layout comes from the torchrun environment (RANK, WORLD_SIZE,
    LOCAL_RANK, MASTER_ADDR, MASTER_PORT); a process started without it is a world of one.
    Replaces ``hvd.init()`` (dist_trainer.py:126)."""
    if _state["initialized"]:
        return
    world = _env_int("WORLD_SIZE", 1)
    if device is None:
        if torch.cuda.is_available() and os.environ.get("GK_FORCE_CPU", "0") != "1":
            device = torch.device("cuda", _env_int("LOCAL_RANK", 0) % max(torch.cuda.device_count(), 1))
        else:
            device = torch.device("cpu")
    if device.type == "cuda":
        torch.cuda.set_device(device)
    else:
        _set_cpu_threads()
    _state["device"] = device
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = backend or ("nccl" if device.type == "cuda" else "gloo")
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = device
        dist.init_process_group(backend=backend, timeout=datetime.timedelta(seconds=timeout_s), **kwargs)
        _state["owns_pg"] = True
    _state["initialized"] = True


def _set_cpu_threads() -> None:
    """CPU workers: split the cores among the workers of this host.  torchrun exports OMP_NUM_THREADS=1 when the user
    set nothing, which leaves all but one core per worker idle (ResNet-20, 2 gloo workers on 8 cores: 52.9 ms/step
    with 1 thread each, 25.3 with 4); mpirun-style launches default to every core in every process, which
    oversubscribes.  ``GK_CPU_THREADS=n`` pins the count; an OMP_NUM_THREADS other than 1 is taken as the user's
    choice and left alone."""
    want = os.environ.get("GK_CPU_THREADS")
    if want:
        n = max(1, int(want))
    else:
        if os.environ.get("OMP_NUM_THREADS", "1") != "1":
            return
        try:
            cores = len(os.sched_getaffinity(0))
        except (AttributeError, OSError):
            cores = os.cpu_count() or 1
        n = max(1, cores // max(_env_int("LOCAL_WORLD_SIZE", _env_int("WORLD_SIZE", 1)), 1))
    try:
        torch.set_num_threads(n)
    except RuntimeError:
        pass


_shutdown_hooks = []


def on_shutdown(fn) -> None:
    """Run ``fn()`` before the process group is destroyed (e.g. drop CUDA graphs that captured collectives:
    destroying a communicator that a live graph still references blocks)."""
    _shutdown_hooks.append(fn)


def shutdown() -> None:
    while _shutdown_hooks:
        try:
            _shutdown_hooks.pop()()
        except Exception:  # noqa: BLE001 - shutting down anyway
            pass
    if _state["owns_pg"] and dist.is_initialized():
        dist.destroy_process_group()
    _state.update(initialized=False, owns_pg=False)


def is_distributed() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def size() -> int:
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def rank() -> int:
    return dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0


def local_rank() -> int:
    return _env_int("LOCAL_RANK", 0)


def local_size() -> int:
    return _env_int("LOCAL_WORLD_SIZE", size())


def device() -> torch.device:
    return _state["device"] or torch.device("cpu")


def _comm_tensor(t: torch.Tensor) -> torch.Tensor:
    """NCCL moves device memory only; gloo host memory only."""
    if not is_distributed():
        return t
    backend = dist.get_backend()
    if backend == "nccl" and not t.is_cuda:
        return t.to(device())
    if backend == "gloo" and t.is_cuda:
        return t.cpu()
    return t


def broadcast(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    """Out-of-place broadcast returning the root's value (``hvd.broadcast``, dist_trainer.py:32-33)."""
    out = tensor.clone()
    if is_distributed():
        buf = _comm_tensor(out)
        dist.broadcast(buf, src=root_rank)
        if buf is not out:
            out.copy_(buf)
    return out


def broadcast_(tensor: torch.Tensor, root_rank: int = 0) -> torch.Tensor:
    if is_distributed():
        buf = _comm_tensor(tensor)
        dist.broadcast(buf, src=root_rank)
        if buf is not tensor:
            tensor.copy_(buf)
    return tensor


def allreduce_(tensor: torch.Tensor, average: bool = True, async_op: bool = False):
    """In-place sum / mean over the workers (``hvd.allreduce_async_``, distributed_optimizer.py:413)."""
    if not is_distributed():
        return None
    if dist.get_backend() == "nccl" and average:
        return dist.all_reduce(tensor, op=dist.ReduceOp.AVG, async_op=async_op)
    work = dist.all_reduce(tensor, op=dist.ReduceOp.SUM, async_op=async_op)
    if average:
        if async_op:
            work.wait()
            work = None
        tensor.div_(size())
    return work


def allgather(tensor: torch.Tensor) -> torch.Tensor:
    """Concatenate equally-shaped tensors in rank order (``hvd.allgather`` for equal counts)."""
    if not is_distributed():
        return tensor.clone()
    out = torch.empty((size() * tensor.shape[0],) + tuple(tensor.shape[1:]), dtype=tensor.dtype,
                      device=tensor.device)
    dist.all_gather_into_tensor(out, tensor.contiguous())
    return out


def barrier() -> None:
    if is_distributed():
        if dist.get_backend() == "nccl":
            dist.barrier(device_ids=[device().index])
        else:
            dist.barrier()


def bcast_object(obj, root: int = 0):
    """Pickled broadcast (replaces mpi4py ``comm.bcast``, dist_trainer.py:40)."""
    if not is_distributed():
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=root)
    return box[0]


def max_over_ranks(value: float) -> float:
    if not is_distributed():
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device() if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
