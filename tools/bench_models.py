#!/usr/bin/env python
"""Throughput of the other configurations BASELINE.json names (bench.py covers the ResNet-50 headline):

    resnet20  CIFAR-10 shape   gaussiank 0.001
    vgg16     CIFAR-10 shape   topk vs gaussiank 0.001
    lstm      PTB shape (2x1500 LSTM, 10k vocabulary)   topk 0.001, gradient clipping 0.25
    lstman4   AN4 spectrogram shape (DeepSpeech)        gaussiank 0.001, clipping 400

Every configuration is driven through the public API exactly like dist_trainer.ssgd():
``trainer.train_and_update(hidden, clip_norm)`` per step, synthetic data of the named shape from the
trainer's own loader (host -> device copy each step), CUDA-event timing with a barrier and synchronize
on both sides, max over ranks.  One JSON line per configuration.

    python tools/bench_models.py [--gpus N] [--steps 30] [--warmup 5] [--only vgg16,lstm] [--out file]
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = [  # name, dnn, dataset, batch per GPU, lr, compressor, density, clip, num_steps
    ("resnet20-gaussiank", "resnet20", "cifar10", 32, 0.1, "gaussiank", 0.001, None, 35),
    ("vgg16-topk", "vgg16", "cifar10", 128, 0.1, "topk", 0.001, None, 35),
    ("vgg16-gaussiank", "vgg16", "cifar10", 128, 0.1, "gaussiank", 0.001, None, 35),
    ("vgg16-dense", "vgg16", "cifar10", 128, 0.1, "none", 1.0, None, 35),
    ("lstm-ptb-topk", "lstm", "ptb", 20, 22.0, "topk", 0.001, 0.25, 35),
    ("lstm-ptb-gaussiank", "lstm", "ptb", 20, 22.0, "gaussiank", 0.001, 0.25, 35),
    ("lstman4-gaussiank", "lstman4", "an4", 4, 0.0002, "gaussiank", 0.001, 400, 35),
]


def run_one(cfg, args):
    import torch
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    from gaussiank_sgd_b200.compression import compressors
    from gaussiank_sgd_b200.dl_trainer import DLTrainer

    name, dnn, dataset, batch, lr, comp, density, clip, num_steps = cfg
    rank, world = hvd.rank(), hvd.size()
    device = hvd.comm.device()
    trainer = DLTrainer(rank, world, dist=False, batch_size=batch, is_weak_scaling=True, ngpus=1,
                        data_dir="synthetic", dataset=dataset, dnn=dnn, lr=lr, nworkers=world, prefix="bench",
                        num_steps=num_steps, synthetic_samples=batch * world * 64)
    opt = hvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                   compression=compressors[comp], is_sparse=density < 1, density=density,
                                   threshold=524288000, force_compress=density < 1)
    hvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    if density < 1:
        opt.increase_one_epoch(); opt.increase_one_epoch()      # density schedule -> the configured value
    hidden = trainer.net.init_hidden() if dnn == "lstm" else None
    for _ in range(max(args.warmup, 3) + 4):
        hidden = trainer.train_and_update(hidden=hidden, clip_norm=clip)
    trainer.flush_metrics()
    torch.cuda.synchronize()
    hvd.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        hidden = trainer.train_and_update(hidden=hidden, clip_norm=clip)
    trainer.flush_metrics()
    e1.record()
    torch.cuda.synchronize()
    hvd.barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t.item())
    if hasattr(opt, "check_errors"):
        opt.check_errors()
    n_params = sum(p.numel() for p in trainer.net.parameters())
    res = {"config": name, "model": dnn, "dataset_shape": dataset, "n_gpus": world, "batch_per_gpu": batch,
           "compressor": comp, "density": density, "clip_norm": clip, "steps": args.steps,
           "ms_per_step": round(ms / args.steps, 4),
           "samples_per_s": round(batch * world * args.steps / (ms * 1e-3), 1),
           "parameters": n_params, "loss": round(float(trainer.loss), 4),
           "cuda_graph": bool(getattr(trainer, "_graph", None) is not None and trainer._graph.graph is not None),
           "own_kernels_per_step": trainer.launches_per_step(), "data": "synthetic", "dtype": "fp32",
           "timing": "CUDA events around K x train_and_update() incl. H2D of inputs, max over ranks"}
    if hasattr(opt, "close"):
        opt.close()
    del trainer, opt
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--only", default=None)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 400),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    hvd.init()
    only = set(args.only.split(",")) if args.only else None
    out = []
    for cfg in CONFIGS:
        if only and cfg[0] not in only and cfg[1] not in only:
            continue
        try:
            res = run_one(cfg, args)
        except Exception as e:  # noqa: BLE001 — one failing configuration must not hide the others
            import traceback
            traceback.print_exc()
            res = {"config": cfg[0], "error": f"{type(e).__name__}: {str(e)[:200]}"}
        if hvd.rank() == 0:
            print(json.dumps(res), flush=True)
            out.append(res)
    if hvd.rank() == 0 and args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump(out, open(args.out, "w"), indent=1)
    hvd.shutdown()


if __name__ == "__main__":
    main()
