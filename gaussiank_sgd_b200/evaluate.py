"""Offline evaluation of saved checkpoints (the reference's evaluate.py:1-75): load
``<dnn>-rank0-epoch<i>.pth`` for i = 1..nepochs, run the trainer's ``test()`` on each and report the best.
Hyper-parameters are parsed from the checkpoint directory name ``<dnn>-n<P>-bs<B>-lr<lr>``."""
from __future__ import annotations

import argparse
import os
import re

import torch

from .dl_trainer import DLTrainer
from .settings import logger


def model_average(trainers):
    """Element-wise mean of several trainers' weights, loaded into the first one (evaluate.py:10-18)."""
    states = [t.net.state_dict() for t in trainers]
    avg = {}
    for k in states[0]:
        if states[0][k].dtype.is_floating_point:
            avg[k] = sum(s[k] for s in states) / len(states)
        else:
            avg[k] = states[0][k].clone()
    trainers[0].net.load_state_dict(avg)
    return trainers[0]


def parse_dirname(path: str):
    name = os.path.basename(os.path.normpath(path))
    m = re.match(r'(?P<dnn>[A-Za-z0-9_]+)-n(?P<n>\d+)-bs(?P<bs>\d+)-lr(?P<lr>[0-9.]+)', name)
    if not m:
        raise ValueError('cannot parse dnn / batch size / lr from %r' % name)
    return m.group('dnn'), int(m.group('n')), int(m.group('bs')), float(m.group('lr'))


def evaluate(model_dir: str, dataset: str, data_dir: str, nepochs: int, ngpus: int = None):
    dnn, nworkers, batch_size, lr = parse_dirname(model_dir)
    ngpus = (1 if torch.cuda.is_available() else 0) if ngpus is None else ngpus
    trainer = DLTrainer(0, 1, dist=False, batch_size=batch_size, is_weak_scaling=True, ngpus=ngpus,
                        data_dir=data_dir, dataset=dataset, dnn=dnn, lr=lr, nworkers=1)
    best_acc, best_epoch, results = None, -1, []
    lower_is_better = dnn in ('lstm', 'lstman4')           # perplexity / WER
    for i in range(1, nepochs + 1):
        fn = os.path.join(model_dir, '%s-rank0-epoch%d.pth' % (dnn, i))
        if not os.path.isfile(fn):
            continue
        trainer.load_model_from_file(fn)
        acc = trainer.test(i)
        results.append((i, acc))
        better = best_acc is None or (acc < best_acc if lower_is_better else acc > best_acc)
        if better:
            best_acc, best_epoch = acc, i
        logger.info('Epoch %d: metric %f (best %s at epoch %d)', i, acc, best_acc, best_epoch)
    logger.info('Best metric: %s at epoch %d', best_acc, best_epoch)
    return best_acc, best_epoch, results


def main(argv=None):
    parser = argparse.ArgumentParser(description='Checkpoint evaluator')
    parser.add_argument('--model-dir', type=str, required=True, help='Directory with <dnn>-rank0-epoch<i>.pth files')
    parser.add_argument('--dataset', type=str, default='cifar10')
    parser.add_argument('--data-dir', type=str, default='./data')
    parser.add_argument('--nepochs', type=int, default=141)
    args = parser.parse_args(argv)
    evaluate(args.model_dir, args.dataset, args.data_dir, args.nepochs)


if __name__ == '__main__':
    main()
