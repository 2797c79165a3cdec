#!/usr/bin/env python
"""Pairs selected per iteration, this framework against the unmodified reference, same recipe.

VERDICT r1 asked whether gaussian-k's under-selection on the merged ResNet-50 buffer (tens to hundreds of
coordinates for k = 25 557 once training is under way) is the algorithm of compression.py:358-389 or an artefact
of this framework's fused path.  Both arms therefore train the same torchvision ResNet-50 from the same seed on
the same synthetic batches, with the reference's density warm-up (0.015625, 0.004, then 0.001; one epoch =
``--epoch-steps`` iterations) and its learning-rate schedule, and log what each selects per iteration:

    torchrun --nproc-per-node 2 tools/selected_trace.py --impl ours      --steps 640 --out gpurun_out/sel_ours.json
    torchrun --nproc-per-node 2 tools/selected_trace.py --impl reference --steps 640 --out gpurun_out/sel_ref.json
    python tools/selected_trace.py --compare gpurun_out/sel_ours.json gpurun_out/sel_ref.json

The reference arm is baseline/_ref (byte-identical to /root/reference) on the torch.distributed stand-ins for
Horovod in baseline/shims, as in bench.py; its count is what it appends to ``_selected_num_gradients``
(distributed_optimizer.py:420).  cuDNN picks its algorithms independently in the two processes, so the two
trajectories are not bit-identical; what is compared is the regime (orders of magnitude of the counts per
density stage) and the loss.
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

BATCH, LR, DENSITY = 32, 0.01, 0.001


def batches(n, rank, pool=64):
    """``pool`` distinct host batches per rank with a class-dependent mean shift (so that the loss can fall),
    cycled; pinned."""
    import torch
    g = torch.Generator().manual_seed(4321 + rank)
    out = []
    for _ in range(pool):
        y = torch.randint(0, 1000, (BATCH,), generator=g)
        x = torch.randn(BATCH, 3, 224, 224, generator=g)
        x += ((y % 16).float() / 8.0 - 1.0).view(-1, 1, 1, 1)
        x[:, 1] += ((y // 16 % 16).float() / 8.0 - 1.0).view(-1, 1, 1)
        out.append((x.pin_memory(), y.pin_memory()))
    return [out[i % pool] for i in range(n)]


def run_ours(args):
    import torch
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    from gaussiank_sgd_b200.compression import compressors
    from gaussiank_sgd_b200.dl_trainer import DLTrainer
    hvd.init()
    rank, world = hvd.rank(), hvd.size()
    torch.manual_seed(1234)
    trainer = DLTrainer(rank, world, dist=False, batch_size=BATCH, is_weak_scaling=True, ngpus=1, data_dir="synthetic",
                        dataset="imagenet", dnn="resnet50", lr=LR, nworkers=world, prefix="seltrace",
                        synthetic_samples=256)
    trainer.num_batches_per_epoch = args.epoch_steps
    opt = hvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                   compression=compressors[args.compressor], is_sparse=True, density=DENSITY,
                                   threshold=524288000, force_compress=True)
    hvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    data = batches(args.steps, rank)
    sel, losses = [], []
    for it in range(args.steps):
        if it and it % args.epoch_steps == 0:
            sel.extend(opt.selected_num_gradients()[-args.epoch_steps:])
            opt.increase_one_epoch()
        trainer.train_and_update(data=data[it])
        trainer.flush_metrics()
        losses.append(round(float(trainer.loss), 5))
    sel.extend(opt.selected_num_gradients()[-(args.steps - len(sel)):])
    torch.cuda.synchronize()
    opt.check_errors()
    res = {"impl": "ours", "selected": sel, "loss": losses, "n_gpus": world, "k_final": int(25557032 * DENSITY)}
    hvd.barrier()
    opt.close()
    hvd.shutdown()
    return res if rank == 0 else None


def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    sys.path[:0] = [os.path.join(ROOT, "baseline", "shims"), ref_dir]
    import logging
    import torch
    import settings as rsettings
    rsettings.LOGGING_GRADIENTS = False
    rsettings.logger.setLevel(logging.WARNING)
    import compression as rcomp
    import distributed_optimizer as rhvd
    import dl_trainer as rdl
    import torch.distributed as dist
    rhvd.init()
    rank, world = rhvd.rank(), rhvd.size()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    torch.manual_seed(1234)
    trainer = rdl.DLTrainer(rank, world, dist=False, batch_size=BATCH, is_weak_scaling=True, ngpus=1, data_dir=None,
                            dataset="imagenet", dnn="resnet50", lr=LR, nworkers=world, prefix="allreduce",
                            pretrain=None, num_steps=35, tb_writer=None)
    trainer.num_batches_per_epoch = args.epoch_steps
    # the epoch roll-over of the reference's train() evaluates on rank 0 and re-seeds the sampler
    # (dl_trainer.py:640-663); without a dataset directory neither attribute exists: one tiny batch, no sampler
    trainer.testloader = [(torch.randn(2, 3, 224, 224), torch.randint(0, 1000, (2,)))]
    trainer.train_sampler = None
    # ... and its test() cannot run on this torch (cal_accuracy views a non-contiguous tensor for top-5,
    # dl_trainer.py:621): the roll-over only evaluates and checkpoints on the trainer whose rank is 0
    trainer.rank = rank + 1000
    key = {"gaussiank": "gaussian"}.get(args.compressor, args.compressor)
    opt = rhvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                    compression=rcomp.compressors[key], is_sparse=True, density=DENSITY,
                                    seq_layernames=None, layerwise_times=None, norm_clip=None, threshold=524288000,
                                    writer=None, gradient_path=None)
    rhvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    data = batches(args.steps, rank)
    sel, losses = [], []
    for it in range(args.steps):
        if it and it % args.epoch_steps == 0:
            sel.extend(opt._selected_num_gradients[-args.epoch_steps:])
            opt.increase_one_epoch()
        opt.zero_grad()
        opt.local = False
        trainer.train(1, data=data[it])
        trainer.update_model()
        losses.append(round(float(trainer.loss), 5))
    sel.extend(opt._selected_num_gradients[-(args.steps - len(sel)):])
    torch.cuda.synchronize()
    res = {"impl": "reference", "selected": [int(s) for s in sel], "loss": losses, "n_gpus": world,
           "k_final": int(25557032 * DENSITY)}
    if dist.is_initialized():
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()
    return res if rank == 0 else None


def compare(a, b, epoch_steps):
    A, B = json.load(open(a)), json.load(open(b))
    n = min(len(A["selected"]), len(B["selected"]))
    rows = []
    for lo in range(0, n, epoch_steps):
        hi = min(lo + epoch_steps, n)
        sa, sb = A["selected"][lo:hi], B["selected"][lo:hi]
        rows.append({"iters": f"{lo}-{hi - 1}", A["impl"] + "_median": statistics.median(sa),
                     B["impl"] + "_median": statistics.median(sb),
                     A["impl"] + "_min_max": [min(sa), max(sa)], B["impl"] + "_min_max": [min(sb), max(sb)],
                     A["impl"] + "_loss": round(sum(A["loss"][lo:hi]) / (hi - lo), 4),
                     B["impl"] + "_loss": round(sum(B["loss"][lo:hi]) / (hi - lo), 4)})
    out = {"n_gpus": A["n_gpus"], "steps": n, "k_final": A["k_final"], "per_stage": rows}
    print(json.dumps(out, indent=1))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--steps", type=int, default=640)
    ap.add_argument("--epoch-steps", type=int, default=64)
    ap.add_argument("--compressor", default="gaussiank")
    ap.add_argument("--out", default=None)
    ap.add_argument("--compare", nargs=2, default=None)
    args = ap.parse_args()
    if args.compare:
        res = compare(args.compare[0], args.compare[1], args.epoch_steps)
        if args.out:
            json.dump(res, open(args.out, "w"), indent=1)
        return
    res = run_reference(args) if args.impl == "reference" else run_ours(args)
    if res is not None and args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump(res, open(args.out, "w"))
        print(args.impl, "selected (every 32nd):", res["selected"][::32], "loss first/last:", res["loss"][0], res["loss"][-1])


if __name__ == "__main__":
    main()
