"""Distributed entry point and S-SGD loop — the reference's dist_trainer.py (ssgd :23-92, CLI :95-136).

Launch: one process per GPU under ``torchrun`` (see run.sh); world layout comes from the environment
instead of ``mpirun``/Horovod.  The flags are the reference's (``--batch-size --nsteps-update --nworkers
--nwpernode --dataset --dnn --data-dir --saved-dir --lr --max-epochs --pretrain --num-steps --compressor
--density --threshold``) plus a few additions documented in ``--help``.
"""
from __future__ import annotations

import argparse
import logging
import os
import time

import numpy as np
import torch

from . import distributed_optimizer as hvd
from . import settings, utils
from .compression import compressors
from .dl_trainer import DLTrainer, _support_datasets, _support_dnns
from .profiling import benchmark
from .settings import formatter, logger

writer = None


NORM_CLIP = {'lstm': 0.25, 'lstman4': 400}      # gradient-norm clipping of the recurrent models (dist_trainer.py:50-54)


def _merge_plan_inputs(trainer, rank):
    """Layer-wise backward profile for the adaptive merge planners (dist_trainer.py:38-46): measured on every
    rank, rank 0's timings are the ones everybody plans with.  (None, None) when the planners are off."""
    if not (settings.ADAPTIVE_MERGE or settings.ADAPTIVE_SPARSE):
        return None, None
    names, times, sizes = benchmark(trainer)
    times = hvd.comm.bcast_object(times, root=0)      # mpi4py comm.bcast in the reference
    if rank == 0:
        logger.info('layerwise backward times: %s', list(times))
        logger.info('layerwise backward sizes: %s', list(sizes))
    logger.info('Benchmarked backward time: %f', np.sum(times))
    logger.info('Model size: %d', np.sum(sizes))
    return names, times


def _resume(trainer, optimizer):
    """Second half of a resume (the first is DLTrainer loading ``--pretrain`` on rank 0): every rank adopts rank
    0's position, optimizer state and schedule stage.  The reference restores the weights only."""
    position = torch.tensor([float(trainer.get_train_epoch()), float(trainer.get_train_iter())])
    position = hvd.broadcast(position, root_rank=0)
    trainer.set_train_epoch(int(position[0]))
    trainer.set_train_iter(int(position[1]))
    hvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(optimizer)
    trainer.restore_optimizer_state()
    optimizer.sync_resume_state(root_rank=0)          # rank 0 alone has read the checkpoint
    optimizer.train_epoch = max(optimizer.train_epoch, trainer.get_train_epoch())
    optimizer._apply_density()


def _step_reference_style(trainer, optimizer, dnn, nsteps_update, norm_clip, hidden):
    """zero_grad / nsteps_update x train(1) / (clip) / update_model — the loop body of dist_trainer.py:70-86, used
    when gradients are accumulated over several mini-batches."""
    optimizer.zero_grad()
    for j in range(nsteps_update):
        optimizer.local = j < nsteps_update - 1       # only the last mini-batch communicates
        if dnn == 'lstm':
            _, hidden = trainer.train(1, hidden=hidden)
        else:
            trainer.train(1)
    if norm_clip is not None:
        optimizer.synchronize()
        optimizer.clip_grad_norm(norm_clip)
    trainer.update_model()
    return hidden


def ssgd(dnn, dataset, data_dir, nworkers, lr, batch_size, nsteps_update, max_epochs, nwpernode, pretrain,
         num_steps, compressor, density, threshold, gradient_path=None, max_iters=None, optimizer_name='sgd',
         synthetic_samples=None, transport=None, dynamic_densities="default", force_compress=False):
    """Synchronous SGD with sparsified gradient exchange (dist_trainer.py:23-92)."""
    rank = hvd.rank()
    use_cuda = torch.cuda.is_available() and os.environ.get("GK_FORCE_CPU", "0") != "1"
    if use_cuda:
        torch.cuda.set_device(hvd.local_rank() % max(torch.cuda.device_count(), 1))
    trainer = DLTrainer(rank, nworkers, dist=False, batch_size=batch_size, is_weak_scaling=True,
                        ngpus=1 if use_cuda else 0, data_dir=data_dir, dataset=dataset, dnn=dnn, lr=lr,
                        nworkers=nworkers, prefix='allreduce', pretrain=pretrain if rank == 0 else None,
                        num_steps=num_steps, tb_writer=writer, optimizer_name=optimizer_name,
                        synthetic_samples=synthetic_samples)
    seq_layernames, layerwise_times = _merge_plan_inputs(trainer, rank)
    norm_clip = NORM_CLIP.get(dnn)
    optimizer = hvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                         compression=compressors[compressor], is_sparse=density < 1, density=density,
                                         seq_layernames=seq_layernames, layerwise_times=layerwise_times,
                                         norm_clip=None, threshold=threshold, writer=writer,
                                         gradient_path=gradient_path, transport=transport,
                                         dynamic_densities=dynamic_densities, force_compress=force_compress)
    _resume(trainer, optimizer)
    iters_per_epoch = max(trainer.get_num_of_training_samples() // (nworkers * batch_size * nsteps_update), 1)
    logger.info('max_epochs: %d', max_epochs)
    display = 40 if iters_per_epoch > 40 else max(iters_per_epoch - 1, 1)
    times, done = [], 0
    for epoch in range(max_epochs):
        hidden = trainer.net.init_hidden() if dnn == 'lstm' else None
        for i in range(iters_per_epoch):
            s = time.time()
            if nsteps_update == 1:
                # zero_grad + forward/backward + exchange + (clip) + update as one (graph-captured) step
                hidden = trainer.train_and_update(hidden=hidden, clip_norm=norm_clip)
            else:
                hidden = _step_reference_style(trainer, optimizer, dnn, nsteps_update, norm_clip, hidden)
            times.append(time.time() - s)
            done += 1
            if i % display == 0 and i > 0:
                if use_cuda:
                    torch.cuda.synchronize()
                time_per_iter = np.mean(times)
                speed = batch_size * nsteps_update / time_per_iter
                logger.warning('Time per iteration including communication: %f, Speed: %f images/s', time_per_iter, speed)
                logger.info('Aggregate speed over %d workers: %f images/s', nworkers, speed * nworkers)
                times = []
            if max_iters is not None and done >= max_iters:
                break
        optimizer.increase_one_epoch()
        if max_iters is not None and done >= max_iters:
            break
    trainer.finish()
    if use_cuda:
        torch.cuda.synchronize()
        optimizer.check_errors()
    return trainer, optimizer


def build_parser():
    parser = argparse.ArgumentParser(description="AllReduce trainer")
    parser.add_argument('--batch-size', type=int, default=32)
    parser.add_argument('--nsteps-update', type=int, default=1)
    parser.add_argument('--nworkers', type=int, default=1, help='Number of workers; under torchrun the world size wins')
    parser.add_argument('--nwpernode', type=int, default=1, help='Number of workers per node')
    parser.add_argument('--dataset', type=str, default='imagenet', choices=_support_datasets, help='Specify the dataset for training')
    parser.add_argument('--dnn', type=str, default='resnet50', choices=_support_dnns, help='Specify the neural network for training')
    parser.add_argument('--data-dir', type=str, default='./data', help='Specify the data root path ("synthetic" forces synthetic data)')
    parser.add_argument('--saved-dir', type=str, default='.', help='Specify the saved weights or gradients root path')
    parser.add_argument('--lr', type=float, default=0.1, help='Default learning rate')
    parser.add_argument('--max-epochs', type=int, default=settings.MAX_EPOCHS, help='Default maximum epochs to train')
    parser.add_argument('--pretrain', type=str, default=None, help='Specify the pretrain path')
    parser.add_argument('--num-steps', type=int, default=35)
    parser.add_argument('--compressor', type=str, default='gaussiank', choices=[k for k in compressors.keys() if k],
                        help='Specify the compressors if \'compression\' is open')
    parser.add_argument('--density', type=float, default=1, help='Density for sparsification')
    parser.add_argument('--threshold', type=int, default=524288000, help='Specify the threshold for gradient merging')
    # additions
    parser.add_argument('--max-iters', type=int, default=None, help='Stop after this many iterations (smoke runs)')
    parser.add_argument('--optimizer', type=str, default='sgd', choices=['sgd', 'lars'])
    parser.add_argument('--synthetic-samples', type=int, default=None, help='Cap the size of synthetic datasets')
    parser.add_argument('--transport', type=str, default=None, choices=['auto', 'p2p', 'symm', 'ipc', 'allgather'])
    parser.add_argument('--no-density-warmup', action='store_true', help='Disable the [0.015625, 0.004, 0.001] schedule')
    parser.add_argument('--force-compress', action='store_true', help='Run the sparsification chain with one worker too')
    return parser


def main(argv=None):
    global writer
    args = build_parser().parse_args(argv)
    hvd.init()
    nworkers = hvd.size() if hvd.size() > 1 else args.nworkers
    if hvd.size() == 1 and args.nworkers > 1:
        logger.warning('--nworkers %d requested but the job has one process; start it with torchrun (run.sh)', args.nworkers)
        nworkers = 1
    batch_size = args.batch_size * args.nsteps_update
    prefix = settings.PREFIX
    if args.density < 1:
        prefix = 'comp-' + args.compressor + '-' + prefix
    logdir = 'allreduce-%s-thres-%dkbytes/%s-n%d-bs%d-lr%.4f-ns%d-ds%s' % (
        prefix, args.threshold / 1024, args.dnn, nworkers, batch_size, args.lr, args.nsteps_update, str(args.density))
    relative_path = './logs/%s' % logdir
    gradient_relative_path = None
    utils.create_path(relative_path)
    if settings.LOGGING_GRADIENTS:
        gradient_relative_path = '%s/gradients/%s' % (args.saved_dir, logdir)
        utils.create_path(gradient_relative_path)
    rank = hvd.rank()
    if rank == 0 and settings.TENSORBOARD:
        try:
            from torch.utils.tensorboard import SummaryWriter
            writer = SummaryWriter('./runs/%s' % logdir)
        except Exception:  # noqa: BLE001 - tensorboard is optional
            writer = None
    logfile = os.path.join(relative_path, settings.hostname + '-' + str(rank) + '.log')
    hdlr = logging.FileHandler(logfile)
    hdlr.setFormatter(formatter)
    logger.addHandler(hdlr)
    logger.info('Configurations: %s', args)
    try:
        ssgd(args.dnn, args.dataset, args.data_dir, nworkers, args.lr, args.batch_size, args.nsteps_update,
             args.max_epochs, args.nwpernode, args.pretrain, args.num_steps, args.compressor, args.density,
             args.threshold, gradient_relative_path, max_iters=args.max_iters, optimizer_name=args.optimizer,
             synthetic_samples=args.synthetic_samples, transport=args.transport,
             dynamic_densities=None if args.no_density_warmup else "default", force_compress=args.force_compress)
    finally:
        hvd.shutdown()


if __name__ == '__main__':
    main()
