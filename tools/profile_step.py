#!/usr/bin/env python
"""Kernel-level breakdown of one training step (torch.profiler over CUDA-graph replays; analysis only, never a
benchmark number).  Prints the top kernels and coarse category sums.

    python tools/profile_step.py [--dnn resnet50] [--batch 32] [--steps 5] [--no-graph] [--out profiles/step.txt]
"""
import argparse
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gaussiank_sgd_b200.distributed_optimizer as hvd  # noqa: E402
from gaussiank_sgd_b200 import settings  # noqa: E402
from gaussiank_sgd_b200.compression import compressors  # noqa: E402
from gaussiank_sgd_b200.dl_trainer import DLTrainer  # noqa: E402

DATASET = {'resnet50': 'imagenet', 'resnet20': 'cifar10', 'vgg16': 'cifar10', 'lstm': 'ptb', 'lstman4': 'an4',
           'googlenet': 'imagenet', 'inceptionv4': 'imagenet', 'alexnet': 'imagenet', 'vgg16i': 'imagenet'}


def category(name: str) -> str:
    n = name.lower()
    if 'k_stem' in n:
        return 'ours (stem conv, tcgen05)'
    if 'k_bn' in n or 'k_pool_gather' in n:
        return 'ours (fused bn+relu)'
    if n.startswith('k_') or 'gkchain' in n:
        return 'ours (gk_*)'
    if 'bn_' in n or 'batch_norm' in n or 'batchnorm' in n:
        return 'batch norm'
    if any(t in n for t in ('conv', 'wgrad', 'dgrad', 'fprop', 'cudnn', 'cutlass', 'gemm', 'xmma', 'sm100', 'sm90', 'nchw', 'nhwc')):
        return 'conv / gemm'
    if 'elementwise' in n or 'vectorized' in n or 'relu' in n or 'add' in n:
        return 'elementwise'
    if 'pool' in n:
        return 'pooling'
    if 'reduce' in n or 'softmax' in n or 'nll' in n:
        return 'reduce / loss'
    if 'memcpy' in n or 'memset' in n:
        return 'memcpy / memset'
    return 'other'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--dnn', default='resnet50')
    ap.add_argument('--batch', type=int, default=32)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--compressor', default='gaussiank')
    ap.add_argument('--out', default=None)
    ap.add_argument('--launches', default=None, help='substring: also list every launch of the matching kernels (us, step 0)')
    a = ap.parse_args()
    settings.CUDA_GRAPH = not a.no_graph
    torch.cuda.set_device(0)
    t = DLTrainer(0, 1, dist=False, batch_size=a.batch, ngpus=1, data_dir='synthetic', dataset=DATASET[a.dnn], dnn=a.dnn,
                  lr=0.01, nworkers=1, synthetic_samples=4096)
    opt = hvd.DistributedOptimizer(t.optimizer, named_parameters=t.net.named_parameters(),
                                   compression=compressors[a.compressor], is_sparse=True, density=0.001,
                                   threshold=524288000, dynamic_densities=None, force_compress=True)
    t.update_optimizer(opt)
    hidden = t.net.init_hidden() if a.dnn == 'lstm' else None
    for _ in range(8):
        hidden = t.train_and_update(hidden=hidden)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(a.steps):
            hidden = t.train_and_update(hidden=hidden)
        torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    per = collections.defaultdict(lambda: [0.0, 0])
    cats = collections.defaultdict(float)
    for e in ev:
        per[e.name][0] += e.device_time / a.steps
        per[e.name][1] += 1
        cats[category(e.name)] += e.device_time / a.steps
    total = sum(v[0] for v in per.values())
    lines = ['%s batch %d graph=%s: %.3f ms of kernel time per step, %d kernels per step' % (
        a.dnn, a.batch, not a.no_graph, total / 1e3, sum(v[1] for v in per.values()) // a.steps)]
    for c, v in sorted(cats.items(), key=lambda kv: -kv[1]):
        lines.append('  %-18s %8.3f ms  %5.1f %%' % (c, v / 1e3, 100 * v / total))
    lines.append('top kernels (us per step, launches per step):')
    for name, (us, cnt) in sorted(per.items(), key=lambda kv: -kv[1][0])[:45]:
        lines.append('  %9.1f  %4d  %s' % (us, cnt // a.steps, name[:150]))
    if a.launches:
        seen = collections.defaultdict(list)
        for e in ev:
            if a.launches in e.name:
                seen[e.name].append(e.device_time)
        lines.append('launches of kernels matching %r (us, all %d profiled steps, in launch order):' % (a.launches, a.steps))
        for name, ts in seen.items():
            n1 = len(ts) // a.steps
            lines.append('  %s' % name[:110])
            lines.append('    ' + ' '.join('%.1f' % x for x in ts[:n1]))
    text = '\n'.join(lines)
    print(text)
    if a.out:
        with open(a.out, 'w') as f:
            f.write(text + '\n')


if __name__ == '__main__':
    main()
