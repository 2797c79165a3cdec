"""Minimal horovod.torch.mpi_ops over torch.distributed (only what the reference calls)."""
import datetime
import os

import torch
import torch.distributed as dist

_handles = {}
_next = [0]


def init():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
            dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=600),
                                    device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
        else:
            dist.init_process_group("gloo", timeout=datetime.timedelta(seconds=600))


def size():
    return dist.get_world_size() if dist.is_initialized() else 1


def rank():
    return dist.get_rank() if dist.is_initialized() else 0


def local_rank():
    return int(os.environ.get("LOCAL_RANK", "0"))


def local_size():
    return int(os.environ.get("LOCAL_WORLD_SIZE", str(size())))


def _new_handle(fn):
    _next[0] += 1
    _handles[_next[0]] = fn
    return _next[0]


def _dev(t):
    if dist.is_initialized() and dist.get_backend() == "nccl" and not t.is_cuda:
        return t.cuda()
    return t


def allreduce_async_(tensor, average=True, name=None):
    if size() == 1:
        return _new_handle(lambda: tensor)
    buf = _dev(tensor)
    work = dist.all_reduce(buf, op=dist.ReduceOp.SUM, async_op=True)

    def finish():
        work.wait()
        if average:
            buf.div_(size())
        if buf is not tensor:
            tensor.copy_(buf)
        return tensor
    return _new_handle(finish)


def allgather_async(tensor, name=None):
    """Variable first-dimension allgather (Horovod semantics): concatenation in rank order."""
    if size() == 1:
        out = tensor.clone()
        return _new_handle(lambda: out)
    t = _dev(tensor.contiguous())
    P = size()
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    sizes = torch.empty(P, dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(sizes, n)
    sizes_l = sizes.tolist()
    m = max(sizes_l)
    pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[:t.shape[0]] = t
    out = torch.empty((P * m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    work = dist.all_gather_into_tensor(out, pad, async_op=True)

    def finish():
        work.wait()
        parts = [out[r * m: r * m + sizes_l[r]] for r in range(P)]
        res = torch.cat(parts, 0)
        return res if tensor.is_cuda else res.cpu()
    return _new_handle(finish)


def broadcast_async_(tensor, root_rank, name=None):
    if size() == 1:
        return _new_handle(lambda: tensor)
    buf = _dev(tensor)
    work = dist.broadcast(buf, src=root_rank, async_op=True)

    def finish():
        work.wait()
        if buf is not tensor:
            tensor.copy_(buf)
        return tensor
    return _new_handle(finish)


def broadcast(tensor, root_rank, name=None):
    out = tensor.clone()
    return synchronize(broadcast_async_(out, root_rank, name))


def synchronize(handle):
    return _handles.pop(handle)()
