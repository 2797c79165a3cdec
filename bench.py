#!/usr/bin/env python
"""Headline benchmark: ResNet-50 (ImageNet shape, 224x224), gaussian-k density 0.001, batch 32 per GPU
(exp_configs/resnet50.conf), fp32, SGD momentum 0.875 — images/s over N GPUs of one node, weak scaling.

    python bench.py --gpus N --steps K --warmup W                       # this framework
    python bench.py --impl reference --gpus N --steps K --warmup W      # the unmodified reference

For N > 1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
--master-port P bench.py --gpus N ...`` (one rank per GPU; RANK/LOCAL_RANK/WORLD_SIZE/MASTER_* from the env).
Rank 0 prints ONE JSON line.

Two timed regions, each bracketed by barrier + torch.cuda.synchronize() and timed with CUDA events on the
launching stream (max over ranks):
  * ``value``  — K steps with the batch already resident on the device (kernel/step time only);
  * ``e2e``    — K steps through the public API (``DLTrainer.train_and_update``): every step copies its inputs
                 from pinned host memory to the device and reads the step's loss back to the host.
The sparsification chain runs in both, also at N = 1 (the reference disables it there,
distributed_optimizer.py:97).  The density warm-up schedule of both arms is advanced to its final value
(0.001) before timing, weights are random-init, data is synthetic.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ResNet-50 gaussiank rho=0.001 images/sec"
BATCH_PER_GPU = 32
LR = 0.01            # exp_configs/resnet50.conf
DENSITY = 0.001


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001 - no nvidia-smi (CPU box)
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); power.append(float(r[3]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def _event_time_ms(fn, device):
    """barrier + sync, CUDA-event timed fn(), sync + barrier; returns max over ranks."""
    import torch
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier(device_ids=[device.index])
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize(device)
    ms = e0.elapsed_time(e1)
    if dist.is_initialized():
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier(device_ids=[device.index])
    return ms


def _host_batches(n, batch, rank, seed=0):
    import torch
    g = torch.Generator().manual_seed(seed + rank)
    out = []
    for _ in range(n):
        x = torch.randn(batch, 3, 224, 224, generator=g).pin_memory()
        y = torch.randint(0, 1000, (batch,), generator=g).pin_memory()
        out.append((x, y))
    return out


def _common_config(world, compressor_key):
    """The config keys both arms print (the driver compares them)."""
    return {"model": "resnet50", "global_batch": BATCH_PER_GPU * world, "seq_len": None, "image": "3x224x224",
            "parallelism": f"dp{world}", "compressor": "gaussiank", "density": DENSITY, "lr": LR,
            "optimizer": "sgd momentum 0.875 wd 6.1e-5", "batch_per_gpu": BATCH_PER_GPU,
            "l2": "no explicit flush: every step streams ~0.4 GB of arenas plus GBs of activations, far larger "
                  "than the 126 MB L2", "registry_key": compressor_key}


def _transport_selfcheck(hvd, compressors, device, world, steps=12):
    """Same seed, same data, two transports: the in-kernel P2P exchange and the NCCL all-gather must leave
    bit-identical parameters on every rank (rank-ordered sums on both paths).  A small MLP (cuBLAS only, no
    atomics in the model) so that any difference can only come from the exchange."""
    import torch
    import torch.nn as nn
    out = {}
    for kind in ("p2p", "allgather"):
        torch.manual_seed(1234)
        model = nn.Sequential(nn.Linear(256, 1024), nn.ReLU(), nn.Linear(1024, 1024), nn.ReLU(),
                              nn.Linear(1024, 100)).to(device)
        base = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
        opt = hvd.DistributedOptimizer(base, named_parameters=model.named_parameters(),
                                       compression=compressors["gaussiank"], is_sparse=True, density=0.01,
                                       threshold=524288000, dynamic_densities=None, transport=kind)
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        g = torch.Generator().manual_seed(77 + hvd.rank())
        for _ in range(steps):
            x = torch.randn(64, 256, generator=g).to(device)
            y = torch.randint(0, 100, (64,), generator=g).to(device)
            opt.zero_grad()
            nn.functional.cross_entropy(model(x), y).backward()
            opt.step()
        torch.cuda.synchronize(device)
        opt.check_errors()
        cs = opt.param_checksum()
        allcs = [None] * world
        torch.distributed.all_gather_object(allcs, cs)
        out[kind] = (allcs, opt._groups[0].transport.name)
        hvd.barrier()
        opt.close()
    same_ranks = all(c == out["p2p"][0][0] for c in out["p2p"][0]) and all(c == out["allgather"][0][0] for c in out["allgather"][0])
    return {"p2p_equals_allgather": bool(same_ranks and out["p2p"][0][0] == out["allgather"][0][0]),
            "selfcheck_replicas_identical": bool(same_ranks), "selfcheck_transports": [out["p2p"][1], out["allgather"][1]],
            "selfcheck_steps": steps}


# --------------------------------------------------------------------------- this framework
def run_ours(args):
    import torch
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    from gaussiank_sgd_b200 import settings
    from gaussiank_sgd_b200.compression import compressors
    from gaussiank_sgd_b200.dl_trainer import DLTrainer
    from gaussiank_sgd_b200.ops import chain as ch
    from gaussiank_sgd_b200.utils.perf_model import measured_peaks

    hvd.init()
    rank, world = hvd.rank(), hvd.size()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE is {world}: launch with torch.distributed.run"
    device = hvd.comm.device()
    assert device.type == "cuda", "bench.py needs a GPU"
    trainer = DLTrainer(rank, world, dist=False, batch_size=BATCH_PER_GPU, is_weak_scaling=True, ngpus=1,
                        data_dir="synthetic", dataset="imagenet", dnn="resnet50", lr=LR, nworkers=world,
                        prefix="bench")
    opt = hvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                   compression=compressors["gaussiank"], is_sparse=True, density=DENSITY,
                                   threshold=524288000, force_compress=True, transport=args.transport)
    hvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    opt.increase_one_epoch(); opt.increase_one_epoch()      # density schedule -> 0.001
    assert opt.get_current_density() == DENSITY
    settings.CUDA_GRAPH = not args.no_graph
    transport = opt._groups[0].transport
    host = trainer.trainloader.batches       # pinned, pre-collated synthetic batches (data/synthetic.py)

    # ---- warm-up (eager steps, graph capture, first replays) through the public API
    for i in range(max(args.warmup, 3) + 4):
        trainer.train_and_update()
    trainer.flush_metrics()
    torch.cuda.synchronize()
    opt.check_errors()
    graphed = trainer._graph is not None and trainer._graph.graph is not None

    # ---- region A: device-resident batch
    dev_batch = tuple(t.to(device) for t in host[0])
    if trainer.channels_last:
        dev_batch = (dev_batch[0].contiguous(memory_format=torch.channels_last), dev_batch[1])
    launches0 = ch.LAUNCHES["n"]

    def region_a():
        for _ in range(args.steps):
            trainer.step_on_device(dev_batch)
    for _ in range(3):
        trainer.step_on_device(dev_batch)
    sampler = ClockSampler(device.index)
    if rank == 0:
        sampler.start()
    ms_a = _event_time_ms(region_a, device)
    per_step_launches = trainer.launches_per_step()
    trainer.flush_metrics()

    # ---- region B: end to end (H2D of the inputs + D2H of the loss every step)
    def region_b():
        for i in range(args.steps):
            trainer.train_and_update()   # next batch from the trainer's loader: pinned host -> device copy
        trainer.flush_metrics()          # the last step's loss is read inside the timed region too
    ms_b = _event_time_ms(region_b, device)
    clocks = sampler.stop() if rank == 0 else None
    opt.check_errors()

    # ---- replica consistency of the run that was just timed: every rank checksums its parameter arena
    cs = opt.param_checksum()
    if world > 1:
        allcs = [None] * world
        torch.distributed.all_gather_object(allcs, cs)
    else:
        allcs = [cs]
    replicas_identical = all(c == allcs[0] for c in allcs)
    phase_means = None
    if args.phases:
        # in-step phase times: eager steps with CUDA events around compress / exchange+update
        settings.CUDA_GRAPH = False
        opt._profiling = True
        opt._phase_events = []
        for _ in range(44):
            trainer.train_and_update()
        trainer.flush_metrics()
        opt._print_profiling()
        phase_means = getattr(opt, "last_phase_means", None)
        opt._profiling = False
        settings.CUDA_GRAPH = not args.no_graph

    # ---- sparse path alone (compress + exchange + receive + zero), eager, CUDA events
    g0 = opt._groups[0]
    sel = opt.selected_num_gradients()
    grad_saved = torch.randn_like(opt._arena.grad) * 1e-3
    resid_saved = opt._resid.clone()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]

    g0.chain.set_grad_sources(None)                    # this micro-measurement reads the gradient arena

    def sparse_only():
        for e0, e1 in ev:
            opt._arena.grad.copy_(grad_saved)      # a fresh gradient for every measured exchange
            opt._resid.copy_(resid_saved)
            e0.record()
            g0.chain.compress(transport.dst())
            transport.exchange()
            ch.scatter(transport.src(), g0.chain.ctrl, opt._arena.grad, 1.0 / world)
            ch.unscatter(transport.src(), g0.chain.ctrl, opt._arena.grad)
            e1.record()
    sparse_only()
    _event_time_ms(sparse_only, device)
    ms_sparse = statistics.median(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        t = torch.tensor([ms_sparse], dtype=torch.float64, device=device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms_sparse = float(t.item())
    n_elem = g0.chain.n_total
    peaks = measured_peaks()
    chain_bytes = 24.0 * n_elem            # K1 16 B/elem (grad+resid read, acc+zero write), K2 4, K3 4
    check = {}
    if world > 1 and not args.no_selfcheck:
        check = _transport_selfcheck(hvd, compressors, device, world)
    result = None
    if rank == 0:
        imgs = BATCH_PER_GPU * world * args.steps
        value = imgs / (ms_a * 1e-3)
        e2e = imgs / (ms_b * 1e-3)
        result = {
            "metric": METRIC, "value": round(value, 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_a / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "gaussiank_sgd_b200",
            "config": _common_config(world, "gaussiank"),
            "details": {"warmup_steps_run": max(args.warmup, 3) + 7,
                        "transport": transport.name + (":" + getattr(transport, "backend", "") if hasattr(transport, "backend") else ""),
                        "cuda_graph": graphed, "channels_last": trainer.channels_last,
                        "folded_update": bool(getattr(trainer._graph, "folded", False)) if graphed else opt.can_fold(),
                        "fused_bn_sites": getattr(trainer, "fused_bn_sites", 0),
                        "model_rewrites": "own BN(+residual)(+ReLU)(+maxpool) kernels at every BatchNorm2d; stem 7x7/2 "
                                          "convolution and its weight gradient by own tcgen05 tf32 kernels on the "
                                          "space-to-depth image (csrc/gk_stem.cu); same parameters, same arithmetic "
                                          "as torchvision resnet50",
                        "stem_tensor_core_calls": dict(__import__("gaussiank_sgd_b200.ops.stem", fromlist=["calls"]).calls),
                        "conv_math": "fp32 storage, cuDNN TF32 convolutions allowed (torch default), fp32 matmul",
                        "phase_means_s": phase_means},
            "replicas_identical": bool(replicas_identical),
            **check,
            "clocks": clocks,
            "e2e": {"value": round(e2e, 2), "unit": "images/s", "ms_per_step": round(ms_b / args.steps, 4),
                    "h2d_bytes_per_step": trainer._prefetch.bytes,
                    "d2h_bytes_per_step": trainer.d2h_bytes},
            "gpu_launches": per_step_launches * args.steps,
            "own_kernels_per_step": per_step_launches,
            "sparse_comm_ms_per_iter": round(ms_sparse, 4),
            "sparse_chain_roofline": {"bytes": chain_bytes, "gbs": round(chain_bytes / (ms_sparse * 1e-3) / 1e9, 1),
                                      "frac_of_" + peaks["source"] + "_hbm": round(chain_bytes / (ms_sparse * 1e-3) / 1e9 / peaks["hbm_gbs"], 3)},
            "selected_per_iter_last": sel[-4:], "k": ch.compute_k(n_elem, DENSITY),
        }
    hvd.barrier()
    opt.close()
    hvd.shutdown()
    return result


# --------------------------------------------------------------------------- reference arm
def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_dir, "distributed_optimizer.py")):
        return {"impl": "reference", "unavailable": "baseline/_ref missing: run tools/install_reference.sh where /root/reference is mounted"}
    sys.path[:0] = [os.path.join(ROOT, "baseline", "shims"), ref_dir]
    for m in ("settings", "utils", "compression", "distributed_optimizer", "dl_trainer", "models", "profiling", "lars"):
        sys.modules.pop(m, None)
    try:
        import torch
        import settings as rsettings
        rsettings.LOGGING_GRADIENTS = False      # the per-iteration gradient dump to disk (Q5) stays off
        import logging
        rsettings.logger.setLevel(logging.WARNING)
        import compression as rcomp
        import distributed_optimizer as rhvd
        import dl_trainer as rdl
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference import failed: {type(e).__name__}: {e}"}
    import torch.distributed as dist
    rhvd.init()
    rank, world = rhvd.rank(), rhvd.size()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE is {world}"
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    trainer = rdl.DLTrainer(rank, world, dist=False, batch_size=BATCH_PER_GPU, is_weak_scaling=True, ngpus=1,
                            data_dir=None, dataset="imagenet", dnn="resnet50", lr=LR, nworkers=world,
                            prefix="allreduce", pretrain=None, num_steps=35, tb_writer=None)
    trainer.num_batches_per_epoch = 1281167 // (BATCH_PER_GPU * world)
    opt = rhvd.DistributedOptimizer(trainer.optimizer, named_parameters=trainer.net.named_parameters(),
                                    compression=rcomp.compressors["gaussian"], is_sparse=True, density=DENSITY,
                                    seq_layernames=None, layerwise_times=None, norm_clip=None, threshold=524288000,
                                    writer=None, gradient_path=None)
    rhvd.broadcast_parameters(trainer.net.state_dict(), root_rank=0)
    trainer.update_optimizer(opt)
    opt.train_epoch = 2                             # density schedule -> 0.001 (what increase_one_epoch does)
    host = _host_batches(4, BATCH_PER_GPU, rank)

    def step(batch):                                # the body of ssgd()'s inner loop (dist_trainer.py:70-86)
        opt.zero_grad()
        opt.local = False
        trainer.train(1, data=batch)
        trainer.update_model()

    for i in range(max(args.warmup, 3) + 4):
        step(host[i % len(host)])
    torch.cuda.synchronize()
    dev_batch = tuple(t.to(device) for t in host[0])
    for _ in range(3):
        step(dev_batch)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_a = _event_time_ms(lambda: [step(dev_batch) for _ in range(args.steps)], device)
    ms_b = _event_time_ms(lambda: [step(host[i % len(host)]) for i in range(args.steps)], device)
    clocks = sampler.stop() if rank == 0 else None
    result = None
    if rank == 0:
        imgs = BATCH_PER_GPU * world * args.steps
        result = {
            "metric": METRIC, "value": round(imgs / (ms_a * 1e-3), 2), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_a / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "impl": "reference",
            "config": _common_config(world, "gaussian"),
            "details": {"warmup_steps_run": max(args.warmup, 3) + 7,
                        "note": "unmodified reference from baseline/_ref driven like ssgd() with train(1, data=batch); "
                                "horovod/mpi4py replaced by torch.distributed stand-ins (baseline/shims); "
                                "settings.LOGGING_GRADIENTS=False; with one worker the reference does not compress "
                                "(distributed_optimizer.py:97)"},
            "clocks": clocks,
            "e2e": {"value": round(imgs / (ms_b * 1e-3), 2), "unit": "images/s", "ms_per_step": round(ms_b / args.steps, 4),
                    "h2d_bytes_per_step": host[0][0].numel() * 4 + host[0][1].numel() * 8, "d2h_bytes_per_step": 4},
            "gpu_launches": 0,
        }
    if dist.is_initialized():
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()
    return result


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--transport", default=None)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-selfcheck", action="store_true", help="skip the P2P == all-gather transport self-check (N > 1)")
    ap.add_argument("--phases", action="store_true", help="also report in-step phase times (44 extra eager steps)")
    args = ap.parse_args()
    if args.gpus > 1 and "RANK" not in os.environ:
        # convenience: re-launch ourselves under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 400), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    # stdout carries exactly one JSON line.  Native libraries write there too (NCCL prints its version banner
    # to file descriptor 1), so descriptor 1 points at stderr while the benchmark runs and the JSON line goes
    # to a private copy of the real stdout.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        res = run_reference(args) if args.impl == "reference" else run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if res is not None and (int(os.environ.get("RANK", "0")) == 0 or "unavailable" not in res):
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
