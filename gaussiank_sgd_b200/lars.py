"""LARS — layer-wise adaptive rate scaling (You, Gitman, Ginsburg 2017) with the arithmetic of the
reference's lars.py:67-134: per-tensor trust ratio eeta*||w|| / (||g|| + wd*||w|| + eps) clamped to
[0, 50], decayed gradient clamped to [-10, 10], an acceleration buffer that starts at ONES, w -= acc.

Used stand-alone it is a plain torch optimizer (per-tensor ops).  Wrapped by
``DistributedOptimizer(..., fused=True)`` the same math runs as two kernels over the arenas
(csrc/gk_cuda.cu k_lars_norms / k_lars_update)."""
import torch
from torch.optim.optimizer import Optimizer


class LARS(Optimizer):
    def __init__(self, params, lr=0.1, momentum=.9, weight_decay=.0005, eeta=0.0001, epsilon=1e-5, max_epoch=200):
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        if eeta < 0.0:
            raise ValueError("Invalid LARS coefficient value: {}".format(eeta))
        defaults = dict(lr=lr, momentum=momentum, weight_decay=weight_decay, eeta=eeta, epsilon=epsilon,
                        max_epoch=max_epoch)
        super().__init__(params, defaults)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            lr, mu, wd = group['lr'], group['momentum'], group['weight_decay']
            eeta, eps = group['eeta'], group['epsilon']
            for p in group['params']:
                if p.grad is None:
                    continue
                w_norm, g_norm = torch.norm(p), torch.norm(p.grad)
                one = torch.ones_like(w_norm)
                trust = torch.where((w_norm > 0) & (g_norm > 0), eeta * w_norm / (g_norm + wd * w_norm + eps), one)
                scaled_lr = lr * trust.clamp_(0.0, 50.0)
                decayed = (p.grad + wd * p).clamp_(-10.0, 10.0)
                state = self.state[p]
                if 'acceleration' not in state:
                    state['acceleration'] = torch.ones_like(p)
                acc = state['acceleration']
                acc.mul_(mu).add_(decayed * scaled_lr)
                p.sub_(acc)
        return loss
