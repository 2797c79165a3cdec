"""Analytic cost models used by the merge planners.

The reference calibrates ``t_topk = s*x*log2(x)`` on a Pascal P102-100 and alpha-beta allgather /
allreduce tables on 1/10 GbE (utils.py:62-134, distributed_optimizer.py:164-169).  The B200
chain is HBM-bound and NVLink latency-bound instead, so the native models are expressed in bytes
moved over the measured copy bandwidth (MEASURED_PEAKS.json) plus a launch/flag latency term.  The
reference's tables are kept (``legacy_*``) so that its planners can be reproduced number for number.
"""
from __future__ import annotations

import json
import math
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def measured_peaks() -> dict:
    p = os.path.join(_ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        d["source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback"}


HBM_GBS = float(measured_peaks().get("hbm_gbs", 6650.0))
NVLINK_GBS = 770.0          # measured peer copy per direction (B200_PROFILING.md)
LAUNCH_S = 4e-6             # one kernel launch + last-block epilogue
FLAG_S = 6e-6               # peer header round trip (release store -> acquire spin)

# bytes per element each chain moves through HBM (csrc/gk_cuda.cu): K1 reads grad+resid, writes
# acc and zeros (16 B); K2 and K3 read acc once each (4 B); top-k adds two digit passes.
CHAIN_BYTES_PER_ELEM = {"gaussian": 24.0, "gaussian2": 24.0, "redsync": 24.0, "redsynctrim": 24.0,
                        "topk": 32.0, "mink": 32.0}
CHAIN_LAUNCHES = {"gaussian": 3, "gaussian2": 3, "redsync": 3, "redsynctrim": 3, "topk": 5, "mink": 5}


def compress_time(x: float, compressor: str = "gaussian") -> float:
    """Seconds of the send-side chain over x elements on one B200."""
    if x <= 0:
        return 0.0
    return CHAIN_LAUNCHES.get(compressor, 3) * LAUNCH_S + x * CHAIN_BYTES_PER_ELEM.get(compressor, 24.0) / (HBM_GBS * 1e9)


def sparse_exchange_time(x: float, P: int, density: float) -> float:
    """Seconds until every peer holds this rank's pairs: (P-1) unicast copies of 8 B/pair."""
    if x <= 0 or P <= 1:
        return 0.0
    return FLAG_S + 8.0 * density * x * (P - 1) / (NVLINK_GBS * 1e9)


def dense_allreduce_time(nbytes: float, P: int) -> float:
    if nbytes <= 0 or P <= 1:
        return 0.0
    return 12e-6 + 2.0 * (P - 1) / P * nbytes / (NVLINK_GBS * 1e9)


# ----------------------------------------------------------------- the reference's own models
LEGACY_TOPK_S = 2.18896957e-10  # utils.py:62 ("P102-100")
LEGACY_GBE_SMALL = {2: (1.6e-3, 1.0e-8), 4: (2.7e-3, 1.3e-8), 8: (4.0e-3, 1.5e-8), 16: (1.7e-3, 1.7e-8)}
LEGACY_GBE_LARGE = {2: (4.4e-3, 5.8e-9), 4: (5.6e-3, 7.4e-9), 8: (7.68e-3, 8.2e-9), 16: (2.1e-3, 1.7e-8)}
LEGACY_10GBE = {2: (1.5e-5, 5.7e-11), 4: (3.6e-5, 1.1e-10), 8: (8.5e-5, 1.4e-10), 16: (1.4e-4, 2.0e-10)}
LEGACY_ALLREDUCE = {16: (0.00010632079996292579, 1.5 * 3.2713239529771973e-10),
                    8: (9.75367204301171e-05, 3.0568230536676206e-10),
                    4: (4.204298980348825e-05, 2.0589360830118177e-10),
                    2: (2.554691138304671e-06, 9.837548167872609e-11)}


def legacy_topk_perf_model(x: float, s: float = LEGACY_TOPK_S) -> float:
    """utils.py:86-93."""
    return 0.0 if x == 0 else s * x * math.log2(x)


def legacy_allgather_perf_model(x: float, P: int, density: float = 0.001) -> float:
    """utils.py:95-108."""
    if x == 0:
        return 0.0
    size = x * P * 4 * density
    a, b = (LEGACY_GBE_LARGE if size >= 1024 * 1024 else LEGACY_GBE_SMALL)[P]
    return (a + b * size) * 2


def legacy_allreduce_time(alpha: float, beta: float, size: float) -> float:
    """utils.py:131-134."""
    return 0.0 if size == 0 else alpha + beta * size


def predict_density_with_size_and_computation(m: float, comp_time: float, P: int) -> float:
    """utils.py:110-129 — the reference computes candidates and then returns 0.001 unconditionally."""
    return 0.001


def get_approximate_sigma_scale(density: float) -> float:
    """utils.py:42-52."""
    if density > 0.7:
        return 0.5
    if density > 0.05:
        return 1.5
    if density > 0.01:
        return 2.0
    return 3.0
