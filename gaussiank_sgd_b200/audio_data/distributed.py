"""Flatten-allreduce data parallelism (the reference carries an apex-era ``DistributedDataParallel`` in
audio_data/distributed.py:17-65 that no entry point uses).  Kept as a small utility: average all gradients
of a module with ONE collective over a flat buffer."""
import torch
import torch.distributed as dist


class DistributedDataParallel(torch.nn.Module):
    def __init__(self, module):
        super().__init__()
        self.module = module
        if dist.is_initialized():
            for p in self.module.state_dict().values():
                if torch.is_tensor(p):
                    dist.broadcast(p, 0)
        self.needs_reduction = False
        for p in self.module.parameters():
            if p.requires_grad:
                p.register_hook(self._mark)

    def _mark(self, grad):
        self.needs_reduction = True
        return grad

    def allreduce_gradients(self):
        """Call after ``backward()``: mean of the gradients over the workers."""
        if not (self.needs_reduction and dist.is_initialized()):
            return
        self.needs_reduction = False
        grads = [p.grad for p in self.module.parameters() if p.grad is not None]
        if not grads:
            return
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat)
        flat /= dist.get_world_size()
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()

    def forward(self, *inputs, **kwargs):
        self.needs_reduction = True
        return self.module(*inputs, **kwargs)
