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
    # merge threshold (elements) as a tenth field: several groups, each launched from the backward hooks on the
    # communication stream as soon as its gradients exist (run with --only resnet50-)
    ("resnet50-gaussiank-merged", "resnet50", "imagenet", 32, 0.01, "gaussiank", 0.001, None, 35, 524288000),
    ("resnet50-gaussiank-groups8M-overlap", "resnet50", "imagenet", 32, 0.01, "gaussiank", 0.001, None, 35, 8388608),
    ("resnet50-gaussiank-groups2M-overlap", "resnet50", "imagenet", 32, 0.01, "gaussiank", 0.001, None, 35, 2097152),
]


def run_one(cfg, args):
    import torch
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    from gaussiank_sgd_b200.compression import compressors
    from gaussiank_sgd_b200.dl_trainer import DLTrainer

    name, dnn, dataset, batch, lr, comp, density, clip, num_steps = cfg[:9]
    threshold = cfg[9] if len(cfg) > 9 else 524288000
    rank, world = hvd.rank(), hvd.size()
    device = hvd.comm.device()
    trainer = DLTrainer(rank, world, dist=False, batch_size=batch, is_weak_scaling=True,
                        ngpus=1 if device.type == "cuda" else 0, data_dir="synthetic", dataset=dataset, dnn=dnn, lr=lr, nworkers=world, prefix="bench",
                        num_steps=num_steps, synthetic_samples=batch * world * 64)
    opt = hvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                   compression=compressors[comp], is_sparse=density < 1, density=density,
                                   threshold=threshold, force_compress=density < 1)
    hvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    if density < 1:
        opt.increase_one_epoch(); opt.increase_one_epoch()      # density schedule -> the configured value
    hidden = trainer.net.init_hidden() if dnn == "lstm" else None
    for _ in range(max(args.warmup, 3) + 4):
        hidden = trainer.train_and_update(hidden=hidden, clip_norm=clip)
    trainer.flush_metrics()
    cuda = device.type == "cuda"
    import time
    if cuda:
        torch.cuda.synchronize()
    hvd.barrier()
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hidden = trainer.train_and_update(hidden=hidden, clip_norm=clip)
    trainer.flush_metrics()
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3            # CPU / gloo run (BASELINE.json config 1)
    hvd.barrier()
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
           "groups": len(getattr(opt, "_groups", [])), "overlap": bool(getattr(opt, "_overlap", False)),
           "timing": "CUDA events around K x train_and_update() incl. H2D of inputs, max over ranks"}
    trainer.finish()
    if hasattr(opt, "close"):
        opt.close()
    del trainer, opt
    import gc
    gc.collect()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
    return res


def _ref_batches(dnn, batch, num_steps, rank, n=4):
    """Synthetic host batches in the formats the reference's train() unpacks (dl_trainer.py:665-680)."""
    import torch
    g = torch.Generator().manual_seed(99 + rank)
    out = []
    for _ in range(n):
        if dnn == "lstm":
            x = torch.randint(0, 10000, (batch, num_steps), generator=g)
            out.append((x, torch.roll(x, -1, 1)))
        elif dnn == "lstman4":
            T = int(torch.randint(100, 400, (1,), generator=g))
            x = torch.randn(batch, 1, 161, T, generator=g)
            sizes = torch.randint(5, 20, (batch,), generator=g).int()
            out.append((x, torch.randint(1, 29, (int(sizes.sum()),), generator=g).int(), torch.ones(batch), sizes))
        else:
            out.append((torch.randn(batch, 3, 32, 32, generator=g), torch.randint(0, 10, (batch,), generator=g)))
    return out


def run_reference_one(cfg, args, mods):
    """The UNMODIFIED reference (baseline/_ref over baseline/shims, as in bench.py --impl reference) on the same
    configuration, driven like the body of its ssgd() loop (dist_trainer.py:70-86).  Its registry key for the
    configured compressor is used as is — at the reference's HEAD ``topk`` runs ``uniform_abs_topk``
    (compression.py:191)."""
    import time
    import torch
    rhvd, rcomp, rdl = mods
    name, dnn, dataset, batch, lr, comp, density, clip, num_steps = cfg
    rank, world = rhvd.rank(), rhvd.size()
    cuda = torch.cuda.is_available() and os.environ.get("GK_FORCE_CPU", "0") != "1"
    if comp == "none" and world > 1:
        return {"config": name, "impl": "reference", "unavailable": "the reference's dense path unpacks three values "
                "from NoneCompressor.compress, which returns two (distributed_optimizer.py:407): it raises with >1 worker"}
    rdl.DLTrainer.vocab_size = 10000          # set by ptb_prepare() only when a data directory exists
    datadir = os.path.join(ROOT, "baseline", "_ref", "audio_data") if dnn == "lstman4" else None
    trainer = rdl.DLTrainer(rank, world, dist=False, batch_size=batch, is_weak_scaling=True, ngpus=1 if cuda else 0,
                            data_dir=None, dataset=dataset, dnn=dnn, lr=lr, nworkers=world, prefix="allreduce",
                            pretrain=None, num_steps=num_steps, tb_writer=None) if dnn != "lstman4" else None
    if trainer is None:
        return {"config": name, "impl": "reference", "unavailable": "DLTrainer('lstman4') needs audio_data/data_loader.py "
                "and decoder.py, which the reference imports and does not ship (dl_trainer.py:456-457)"}
    trainer.num_batches_per_epoch = 10 ** 6
    key = {"gaussiank": "gaussian"}.get(comp, comp)
    opt = rhvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                    compression=rcomp.compressors[key], is_sparse=density < 1, density=density,
                                    seq_layernames=None, layerwise_times=None, norm_clip=None, threshold=524288000,
                                    writer=None, gradient_path=None)
    rhvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    opt.train_epoch = 2                        # density schedule -> the configured value
    data = _ref_batches(dnn, batch, num_steps, rank)
    hidden = [trainer.net.init_hidden()] if dnn == "lstm" else [None]

    def step(i):
        opt.zero_grad()
        opt.local = False
        if dnn == "lstm":
            _, hidden[0] = trainer.train(1, data=data[i % len(data)], hidden=hidden[0])
        else:
            trainer.train(1, data=data[i % len(data)])
        if clip is not None:
            opt.synchronize()
            torch.nn.utils.clip_grad_norm_(trainer.net.parameters(), clip)
        trainer.update_model()

    for i in range(max(args.warmup, 3) + 4):
        step(i)
    if cuda:
        torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    for i in range(args.steps):
        step(i)
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda" if cuda else "cpu")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t.item())
    res = {"config": name, "impl": "reference", "model": dnn, "n_gpus": world if cuda else 0, "n_workers": world,
           "batch_per_gpu": batch, "compressor": key, "density": density, "clip_norm": clip, "steps": args.steps,
           "ms_per_step": round(ms / args.steps, 4), "samples_per_s": round(batch * world * args.steps / (ms * 1e-3), 1),
           "data": "synthetic", "dtype": "fp32",
           "note": "one worker: the reference does not compress (distributed_optimizer.py:97)" if world == 1 else ""}
    del trainer, opt
    if cuda:
        torch.cuda.empty_cache()
    return res


def _reference_modules():
    sys.path[:0] = [os.path.join(ROOT, "baseline", "shims"), os.path.join(ROOT, "baseline", "_ref")]
    import logging
    import settings as rsettings
    rsettings.LOGGING_GRADIENTS = False
    rsettings.logger.setLevel(logging.WARNING)
    import compression as rcomp
    import distributed_optimizer as rhvd
    import dl_trainer as rdl
    rhvd.init()
    return rhvd, rcomp, rdl


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--only", default=None)
    ap.add_argument("--out", default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 400),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.impl == "reference":
        if not os.path.isfile(os.path.join(ROOT, "baseline", "_ref", "distributed_optimizer.py")):
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref missing: run tools/install_reference.sh"}))
            return
        mods = _reference_modules()
        hvd = mods[0]
        hvd.barrier = lambda: None
        hvd.shutdown = lambda: None
    else:
        import gaussiank_sgd_b200.distributed_optimizer as hvd
        hvd.init()
    only = set(args.only.split(",")) if args.only else None
    out = []
    for cfg in CONFIGS:
        if only is None and len(cfg) > 9:
            continue                                   # the grouping study only runs when asked for
        if only and cfg[0] not in only and cfg[1] not in only:
            continue
        try:
            res = run_reference_one(cfg, args, mods) if args.impl == "reference" else run_one(cfg, args)
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
