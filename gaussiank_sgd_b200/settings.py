"""Global switches and the shared logger.

The reference keeps these as module constants that are edited by hand (settings.py:1-44).  The
names and defaults are preserved, but every switch can be set from the environment (``GK_<NAME>``),
and ``LOGGING_GRADIENTS`` is off by default: the reference copies the whole gradient to the host
and ``np.save``s it on rank 0 every iteration (distributed_optimizer.py:422-424).
"""
from __future__ import annotations

import logging
import os
import socket


def _flag(name: str, default: bool) -> bool:
    v = os.environ.get("GK_" + name)
    if v is None:
        return default
    return v.strip().lower() not in ("0", "false", "no", "off", "")


def _floats(name: str, default):
    v = os.environ.get("GK_" + name)
    if v is None:
        return default
    v = v.strip()
    if v.lower() in ("", "none", "off"):
        return None
    return [float(x) for x in v.split(",")]


DEBUG = _flag("DEBUG", False)

SERVER_PORT = 5911  # vestiges of the reference's parent project, unused there as well
PORT = 5922

WARMUP = _flag("WARMUP", True)

PREFIX = ''
if WARMUP:
    PREFIX = PREFIX + 'gwarmup'

LOGGING_ASSUMPTION = _flag("LOGGING_ASSUMPTION", False)
LOGGING_GRADIENTS = _flag("LOGGING_GRADIENTS", False)
GRADIENT_DUMP_EVERY = int(os.environ.get("GK_GRADIENT_DUMP_EVERY", "1"))

EXP = os.environ.get("GK_EXP", '-convergence')
PREFIX = PREFIX + EXP
ADAPTIVE_MERGE = _flag("ADAPTIVE_MERGE", False)
ADAPTIVE_SPARSE = _flag("ADAPTIVE_SPARSE", False)
if ADAPTIVE_MERGE:
    PREFIX = PREFIX + '-ada'

TENSORBOARD = _flag("TENSORBOARD", False)
USE_FP16 = _flag("USE_FP16", False)                  # mixed precision (settings.py:28): autocast forward/backward,
#   fp32 master weights / gradients / residuals / exchange in the arenas (what apex O2 keeps in fp32 too)
AMP_DTYPE = os.environ.get("GK_AMP_DTYPE", "fp16")   # fp16 (the reference's apex O2, static loss scale 128) | bf16
AMP_LOSS_SCALE = float(os.environ.get("GK_AMP_LOSS_SCALE", "128"))   # dl_trainer.py:41-45

MAX_EPOCHS = int(os.environ.get("GK_MAX_EPOCHS", "10"))

# ---- switches that exist only here -------------------------------------------------------------
DYNAMIC_DENSITIES = _floats("DYNAMIC_DENSITIES", [0.015625, 0.004, 0.001])  # distributed_optimizer.py:60
COMPRESS_SINGLE = _flag("COMPRESS_SINGLE", False)   # run the chain with one worker too
STEAL_GRADS = _flag("STEAL_GRADS", True)            # K1 reads autograd's gradient tensors in place (no arena accumulate)
SPARSE_ZERO = _flag("SPARSE_ZERO", True)            # zero_grad clears only the received coordinates
FOLDED_UPDATE = _flag("FOLDED_UPDATE", True)        # step(): receive folded into the SGD kernel (no dense gradient)
PROFILE_PHASES = _flag("PROFILE_PHASES", False)     # CUDA-event timers for compress / comm / update
NVTX = _flag("NVTX", False)                          # NVTX ranges around forward / backward / compress / comm / update
VERIFY_EVERY = int(os.environ.get("GK_VERIFY_EVERY", "0"))   # >0: cross-check the fused path against the oracle every N steps
CUDA_GRAPH = _flag("CUDA_GRAPH", True)              # capture the training step in a CUDA graph on GPUs
NCCL_GRAPH = _flag("NCCL_GRAPH", True)              # also capture steps that contain NCCL collectives (dense / all-gather paths)
CHANNELS_LAST = _flag("CHANNELS_LAST", True)        # NHWC convolutions on GPUs
SYNTHETIC_POOL = _flag("SYNTHETIC_POOL", True)      # pre-collated pinned batches for synthetic image data
FUSED_BN = _flag("FUSED_BN", True)                  # BN(+residual)+ReLU through csrc/gk_bn.cu on GPUs (ops/bn.py)
FUSED_BN_COOP = _flag("FUSED_BN_COOP", True)        # single cooperative kernel per direction where the shape allows
FUSED_BN_POOL = _flag("FUSED_BN_POOL", True)        # stem: BN+ReLU+maxpool(3,2,1) without the full-resolution write
FUSED_BN_MASK = _flag("FUSED_BN_MASK", True)        # residual junctions: ReLU bit mask for the backward instead of y
FUSED_BN_PAIR = _flag("FUSED_BN_PAIR", True)        # downsample blocks: bn3 and the shortcut's BN in one kernel
FUSED_BN_TWICE = _flag("FUSED_BN_TWICE", True)      # block outputs handed out twice: no autograd add at the junctions
STEM_S2D = _flag("STEM_S2D", True)                  # ResNet stem as a 4x4 convolution on the space-to-depth image
STEM_TC = _flag("STEM_TC", True)                    # ... and that 4x4 form on tcgen05 tensor cores (csrc/gk_stem.cu)
STEM_TC_WGRAD = _flag("STEM_TC_WGRAD", True)        # ... and its weight gradient (k_stem_wgrad_tc) instead of cuDNN's
STEM_BN_STATS = _flag("STEM_BN_STATS", True)        # ... its epilogue accumulates the batch statistics of bn1
STEM_PAD4 = _flag("STEM_PAD4", True)                # ResNet stem: pad the 3 input channels to 4 (aligned cuDNN kernels)
BN_STAMPS = _flag("BN_STAMPS", False)                # fused BN kernels record phase time stamps (profiling aid)
# cuDNN autotuning.  Unset = per workload (DLTrainer): off for the ImageNet-shape models, on for the others.  Measured on
# ResNet-50: on one GPU the heuristic choice is as fast as the autotuned one (5 115 against 5 071-5 117 images/s), on 8
# GPUs it is 1.7 % faster (6.48 against 6.59 ms/step) because every rank then runs the same algorithms - the autotuner
# times each candidate once, ranks land on different kernels, and the step runs at the pace of the slowest rank.
# (The reference pins benchmark=False, deterministic=True.)
CUDNN_BENCHMARK = (_flag("CUDNN_BENCHMARK", True) if os.environ.get("GK_CUDNN_BENCHMARK") is not None else None)

CUDNN_BENCHMARK_LIMIT = (int(os.environ["GK_CUDNN_BENCHMARK_LIMIT"])
                         if os.environ.get("GK_CUDNN_BENCHMARK_LIMIT") else None)   # None = torch's default (10)

hostname = socket.gethostname()
logger = logging.getLogger(hostname)

if DEBUG:
    logger.setLevel(logging.DEBUG)
else:
    logger.setLevel(logging.INFO)

formatter = logging.Formatter('%(asctime)s [%(filename)s:%(lineno)d] %(levelname)s %(message)s')
if not logger.handlers:
    strhdlr = logging.StreamHandler()
    strhdlr.setFormatter(formatter)
    logger.addHandler(strhdlr)
logger.propagate = False
