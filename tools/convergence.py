#!/usr/bin/env python
"""Convergence parity on synthetic data: the same model, initial weights, data order and learning-rate
schedule trained with dense gradients, exact top-k and gaussian-k (error feedback on), loss logged per step.

The north star asks that gaussian-k "match top-k convergence on synthetic data of the named shape with
random-init weights" (the paper's claim, README of the reference: gaussian-k reaches top-k accuracy).
The synthetic images carry a class-dependent mean shift (data/synthetic.py), so the loss can fall.

    python tools/convergence.py --dnn resnet50 --steps 300 [--gpus 2] [--density 0.001] [--out profiles/x.json]

With one GPU the compressors still run (force_compress): selection, residual and update are exercised,
the exchange is the local transport.  With N GPUs every rank sees its own shard of the pool.
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = {"resnet50": ("imagenet", 32, 0.01), "resnet20": ("cifar10", 32, 0.1), "vgg16": ("cifar10", 128, 0.1),
          "lstm": ("ptb", 20, 22.0)}


def run(comp, density, args):
    import torch
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    from gaussiank_sgd_b200.compression import compressors
    from gaussiank_sgd_b200.dl_trainer import DLTrainer

    dataset, batch, lr = SHAPES[args.dnn]
    if args.batch:
        batch = args.batch
    rank, world = hvd.rank(), hvd.size()
    torch.manual_seed(1234)                      # identical initial weights for every compressor
    trainer = DLTrainer(rank, world, dist=False, batch_size=batch, is_weak_scaling=True, ngpus=1,
                        data_dir="synthetic", dataset=dataset, dnn=args.dnn, lr=args.lr or lr, nworkers=world,
                        prefix="conv", num_steps=35, synthetic_samples=args.samples)
    opt = hvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                   compression=compressors[comp], is_sparse=density < 1, density=density,
                                   threshold=args.threshold, force_compress=density < 1,
                                   # "default": the reference's warm-up 0.015625, 0.004, then the target
                                   # (one value per epoch of the synthetic pool); None: the target from step 0
                                   dynamic_densities="default" if args.schedule == "default" else None)
    hvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    clip = 0.25 if args.dnn == "lstm" else None
    hidden = trainer.net.init_hidden() if args.dnn == "lstm" else None
    losses = []
    per_epoch = max(trainer.get_num_of_training_samples() // (world * batch), 1)
    for it in range(args.steps):
        if it and it % per_epoch == 0 and hasattr(opt, "increase_one_epoch"):
            opt.increase_one_epoch()             # what ssgd() does after every epoch: advances the density schedule
        hidden = trainer.train_and_update(hidden=hidden, clip_norm=clip)
        trainer.flush_metrics()
        losses.append(round(float(trainer.loss), 5))
    if world > 1:                                # mean loss over ranks
        t = torch.tensor(losses, dtype=torch.float64, device=hvd.comm.device())
        torch.distributed.all_reduce(t)
        losses = [round(v / world, 5) for v in t.tolist()]
    sel = opt.selected_num_gradients()[-5:] if density < 1 and hasattr(opt, "selected_num_gradients") else None
    if hasattr(opt, "check_errors"):
        opt.check_errors()
    if hasattr(opt, "close"):
        opt.close()
    del trainer, opt
    torch.cuda.empty_cache() if torch.cuda.is_available() else None
    return losses, sel


def smooth(v, w):
    out = []
    for i in range(len(v)):
        seg = v[max(0, i - w + 1):i + 1]
        out.append(sum(seg) / len(seg))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dnn", default="resnet50", choices=sorted(SHAPES))
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--density", type=float, default=0.001)
    ap.add_argument("--samples", type=int, default=2048, help="size of the synthetic training pool")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--lr", type=float, default=None)
    ap.add_argument("--compressors", default="none,topk,gaussiank")
    ap.add_argument("--threshold", type=int, default=524288000,
                    help="gradient merging threshold in bytes: 524288000 = one merged buffer (the reference's "
                         "default), 0 = every tensor is sparsified on its own")
    ap.add_argument("--schedule", default="none", choices=["none", "default"],
                    help="density warm-up over the first epochs (the reference's [0.015625, 0.004, 0.001]) or none")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 400),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    hvd.init()
    curves, selected = {}, {}
    for comp in args.compressors.split(","):
        d = 1.0 if comp == "none" else args.density
        curves[comp], selected[comp] = run(comp, d, args)
    if hvd.rank() == 0:
        w = max(args.steps // 10, 1)
        tail = {c: round(sum(v[-w:]) / w, 4) for c, v in curves.items()}
        res = {"dnn": args.dnn, "n_gpus": hvd.size(), "steps": args.steps, "density": args.density,
               "merge_threshold_bytes": args.threshold, "density_schedule": args.schedule,
               "pool": args.samples, "first_loss": {c: v[0] for c, v in curves.items()},
               "mean_loss_last_10pct": tail, "selected_per_iter_last": selected,
               "smoothed_every_10pct": {c: [round(x, 4) for x in smooth(v, w)[w - 1::w]] for c, v in curves.items()},
               "curves": curves}
        print(json.dumps({k: v for k, v in res.items() if k != "curves"}), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
            json.dump(res, open(args.out, "w"))
    hvd.shutdown()


if __name__ == "__main__":
    main()
