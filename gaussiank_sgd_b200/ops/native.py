"""ctypes bindings of the two native libraries (include/gk_b200.h).

``libgk_cuda.so`` (csrc/gk_cuda.cu, sm_100a) and ``libgk_cpu.so`` (csrc/gk_cpu.cpp) are built
IN-TREE next to this file so that they travel with the repository snapshot to the GPU box.
The structures below mirror the C header field by field.

The reference has no native code; every kernel here replaces an implicit
ATen/Horovod kernel listed in SURVEY.md §2.5.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
CSRC = os.path.join(_ROOT, "csrc")
INCLUDE = os.path.join(_ROOT, "include")
# GK_CUDA_SO points at an alternative build (tuning sweeps: same sources, other -D flags)
CUDA_SO = os.environ.get("GK_CUDA_SO") or os.path.join(_HERE, "libgk_cuda.so")
CPU_SO = os.path.join(_HERE, "libgk_cpu.so")

GK_NC = 16
GK_MAXP = 16
GK_HIST_BINS = 2048
GK_THREADS = 256
GK_SLOT_HDR = 2

MODE_GAUSSIAN, MODE_TOPK, MODE_REDSYNC, MODE_REDSYNCTRIM, MODE_THRESHOLD, MODE_MINK, MODE_DGC = range(7)


class GkBlk(C.Structure):
    _fields_ = [("start", C.c_int32), ("len", C.c_int32), ("seg", C.c_int32), ("soff", C.c_int32)]


class GkSeg(C.Structure):
    _fields_ = [("n", C.c_int32), ("k", C.c_int32), ("blk_begin", C.c_int32), ("blk_end", C.c_int32),
                ("zabs", C.c_double), ("param", C.c_float), ("aux", C.c_int32)]


class GkSegStat(C.Structure):
    _fields_ = [("mean", C.c_float), ("std", C.c_float), ("meanabs", C.c_float), ("maxabs", C.c_float),
                ("cand", C.c_float * GK_NC), ("thr", C.c_float), ("tbits", C.c_uint32), ("ties", C.c_uint32),
                ("chosen", C.c_int32), ("count", C.c_uint32), ("out_base", C.c_uint32),
                ("prefix", C.c_uint32), ("krem", C.c_uint32), ("rcand", C.c_uint8 * GK_NC), ("nrc", C.c_int32),
                ("tot", C.c_uint32 * GK_NC), ("known", C.c_uint32), ("gate", C.c_int32), ("done", C.c_int32),
                ("sprefix", C.c_uint32), ("skrem", C.c_uint32)]


class GkCtrl(C.Structure):
    _fields_ = [("done", C.c_uint32 * 8), ("total", C.c_uint32), ("sent", C.c_uint32), ("seq", C.c_uint32),
                ("err", C.c_uint32), ("overflow", C.c_uint32), ("rsv", C.c_uint32 * 3),
                ("sent_total", C.c_uint64), ("ring", C.c_uint32 * 64)]


class GkChain(C.Structure):
    _fields_ = [("grad", C.c_void_p), ("resid", C.c_void_p), ("blk", C.c_void_p), ("seg", C.c_void_p),
                ("part", C.c_void_p), ("stat", C.c_void_p), ("bcnt", C.c_void_p), ("boff", C.c_void_p),
                ("hist", C.c_void_p), ("ctrl", C.c_void_p),
                ("nblk", C.c_int32), ("nseg", C.c_int32), ("mode", C.c_int32), ("nloops", C.c_int32),
                ("use_residual", C.c_int32), ("zero_grad", C.c_int32), ("cap", C.c_uint32), ("rsv", C.c_int32),
                ("gsrc", C.c_void_p), ("rowmax", C.c_void_p), ("rpb", C.c_int32), ("rsv2", C.c_int32)]


class GkDst(C.Structure):
    _fields_ = [("slot", C.c_void_p * GK_MAXP), ("ndst", C.c_int32), ("rsv", C.c_int32),
                ("parity_stride", C.c_uint64)]


class GkSrc(C.Structure):
    _fields_ = [("arena", C.c_void_p), ("slot_stride", C.c_uint64), ("parity_stride", C.c_uint64),
                ("nsrc", C.c_int32), ("wait", C.c_int32), ("timeout_ns", C.c_uint64)]


class GkHyper(C.Structure):
    _fields_ = [("lr", C.c_float), ("momentum", C.c_float), ("dampening", C.c_float),
                ("weight_decay", C.c_float), ("nesterov", C.c_int32), ("first_step", C.c_int32),
                ("lars_eeta", C.c_float), ("lars_eps", C.c_float)]


class GkRecvUpd(C.Structure):
    _fields_ = [("param", C.c_void_p), ("mom", C.c_void_p), ("mask", C.c_void_p), ("blk", C.c_void_p),
                ("hyper", C.c_void_p), ("gmap", C.c_void_p), ("ctrl", C.c_void_p), ("bar", C.c_void_p),
                ("scale", C.c_float), ("nblk", C.c_int32), ("nelem", C.c_int32), ("skip_dense", C.c_int32)]


assert C.sizeof(GkBlk) == 16 and C.sizeof(GkSeg) == 32 and C.sizeof(GkSegStat) == 216
assert C.sizeof(GkCtrl) == 328 and C.sizeof(GkHyper) == 32

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]
# AVX2 + FMA (any x86-64 host of a B200 has them): the dense loops vectorise and std::fmaf is one instruction;
# -ffp-contract=off keeps every other multiply-add unfused, as in the CUDA twins
GXX_FLAGS = ["-O3", "-mavx2", "-mfma", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC"]


def _stale(so: str, srcs) -> bool:
    if not os.path.exists(so):
        return True
    t = os.path.getmtime(so)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in srcs)


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


CUDA_SOURCES = ("gk_cuda.cu", "gk_bn.cu", "gk_stem.cu")


def _cuda_deps():
    return [os.path.join(CSRC, f) for f in CUDA_SOURCES] + [os.path.join(INCLUDE, "gk_b200.h")]


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    """Cross-compile csrc/gk_cuda.cu + csrc/gk_bn.cu + csrc/gk_stem.cu for sm_100a into one library (works without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in CUDA_SOURCES]
    if force or _stale(CUDA_SO, _cuda_deps()):
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", CUDA_SO + ".tmp"] + srcs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
        os.replace(CUDA_SO + ".tmp", CUDA_SO)
        if verbose:
            print(r.stderr)
    return CUDA_SO


def _host_has_avx2_fma() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = set(line.split(":", 1)[1].split())
                    return "avx2" in flags and "fma" in flags
    except OSError:
        pass
    return True          # no /proc/cpuinfo (not Linux / x86): leave the decision to the compiler flags below


def build_cpu(force: bool = False) -> str:
    """libgk_cpu.so (AVX2 + FMA).  A host without those extensions - the in-tree library travels between machines -
    gets its own plain build next to it."""
    src = os.path.join(CSRC, "gk_cpu.cpp")
    hdr = os.path.join(INCLUDE, "gk_b200.h")
    so, flags = CPU_SO, GXX_FLAGS
    if not _host_has_avx2_fma():
        so, flags = CPU_SO.replace(".so", "_generic.so"), [f for f in GXX_FLAGS if f not in ("-mavx2", "-mfma")]
    if force or _stale(so, [src, hdr]):
        cmd = ["g++"] + flags + ["-o", so + ".tmp", src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
        os.replace(so + ".tmp", so)
    return so


_lock = threading.Lock()
_libs = {}

_P = C.c_void_p


def _proto(lib, name, argtypes, restype=C.c_int):
    f = getattr(lib, name)
    f.argtypes = argtypes
    f.restype = restype
    return f


def cuda_lib():
    """Load (building if needed) the sm_100a library.  Fails loudly: on a GPU box the native path
    is the only path."""
    with _lock:
        if "cuda" not in _libs:
            if _stale(CUDA_SO, _cuda_deps()):
                build_cuda()
            lib = C.CDLL(CUDA_SO)
            ch, ds, sr = C.POINTER(GkChain), C.POINTER(GkDst), C.POINTER(GkSrc)
            _proto(lib, "gk_version", [], C.c_char_p)
            for n in ("gk_accumulate", "gk_count_pick", "gk_topk_select"):
                _proto(lib, n, [ch, _P])
            _proto(lib, "gk_compact_send", [ch, ds, _P])
            _proto(lib, "gk_compress", [ch, ds, _P])
            _proto(lib, "gk_dgc_select", [ch, C.c_uint64, _P])
            _proto(lib, "gk_dgc_compress", [ch, ds, C.c_uint64, _P])
            _proto(lib, "gk_scatter", [sr, _P, _P, C.c_float, C.c_int32, _P])
            _proto(lib, "gk_unscatter", [sr, _P, _P, C.c_int32, _P])
            _proto(lib, "gk_sgd", [_P, _P, _P, _P, C.c_int32, _P, _P, _P])
            _proto(lib, "gk_sgd_recv", [C.POINTER(GkRecvUpd), sr, _P])
            _proto(lib, "gk_lars", [_P, _P, _P, _P, C.c_int32, _P, C.c_int32, _P, _P, _P, _P, _P])
            _proto(lib, "gk_grad_norm", [_P, _P, C.c_int32, _P, C.c_float, _P, _P, _P])
            _proto(lib, "gk_stem_fwd", [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P])
            _proto(lib, "gk_stem_fwd_stats", [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P])
            _proto(lib, "gk_stem_stats_rows", [])
            _proto(lib, "gk_bn_pool_fwd_sums", [_P] * 11 + [C.c_int32] * 5 + [C.c_float, C.c_float, _P])
            _proto(lib, "gk_stem_s2d", [_P, _P, C.c_int32, C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_stem_s2d_layout", [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_stem_pack_w", [_P, _P, _P])
            _proto(lib, "gk_stem_supported", [C.c_int32, C.c_int32])
            _proto(lib, "gk_stem_driver_ok", [_P])
            _proto(lib, "gk_stem_wgrad_supported", [C.c_int32, C.c_int32])
            _proto(lib, "gk_stem_wgrad_workspace", [], C.c_longlong)
            _proto(lib, "gk_stem_wgrad", [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P])
            _proto(lib, "gk_grad_norm_static", [_P, _P, C.c_int32, C.c_float, _P, _P])
            _proto(lib, "gk_scale", [_P, _P, C.c_int32, _P, _P])
            _proto(lib, "gk_randomk_send", [ch, ds, C.c_uint64, _P])
            _proto(lib, "gk_bucket_means", [ch, _P, _P])
            _proto(lib, "gk_bucket_reconstruct", [ch, _P, _P, _P])
            _proto(lib, "gk_copy", [_P, _P, C.c_int64, _P])
            _proto(lib, "gk_histogram", [_P, C.c_int64, C.c_float, C.c_float, C.c_int32, _P, _P])
            _proto(lib, "gk_bn_fwd", [_P] * 12 + [C.c_int64, C.c_int32, C.c_float, C.c_float, C.c_int32, _P])
            _proto(lib, "gk_bn_bwd", [_P] * 14 + [C.c_int64, C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_bn_fused_supported", [C.c_int64, C.c_int32])
            _proto(lib, "gk_bn_set_tuning", [C.c_int32, C.c_int32])
            _proto(lib, "gk_bn_set_big_mb", [C.c_int32])
            _proto(lib, "gk_bn_set_pair_big_mb", [C.c_int32])
            for env, fn in (("GK_BN_BIG_MB", lib.gk_bn_set_big_mb), ("GK_BN_PAIR_BIG_MB", lib.gk_bn_set_pair_big_mb)):
                if os.environ.get(env):          # tuning knobs of the fused BN geometry (include/gk_b200.h)
                    check(fn(int(os.environ[env])), env)
            _proto(lib, "gk_bn_bwd_fused2", [_P] * 14 + [C.c_int64, C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_bn2_fwd_fused", [_P, _P, C.c_float, C.c_float, C.c_float, C.c_float, _P, _P, _P, _P, C.c_int64,
                                             C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_bn_fwd_fused_m", [_P] * 12 + [C.c_int64, C.c_int32, C.c_float, C.c_float, C.c_int32, _P, _P])
            _proto(lib, "gk_bn2_bwd_fused", [_P, _P, _P, _P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_bn_pool_fwd", [_P] * 13 + [C.c_int32] * 4 + [C.c_float, C.c_float, _P])
            _proto(lib, "gk_bn_pool_bwd", [_P] * 13 + [C.c_int32] * 5 + [_P])
            _proto(lib, "gk_bn_fwd_fused", [_P] * 12 + [C.c_int64, C.c_int32, C.c_float, C.c_float, C.c_int32, _P])
            _proto(lib, "gk_bn_bwd_fused", [_P] * 13 + [C.c_int64, C.c_int32, C.c_int32, _P])
            _proto(lib, "gk_ipc_alloc", [C.c_uint64, C.POINTER(C.c_void_p), C.c_char_p])
            _proto(lib, "gk_ipc_open", [C.c_char_p, C.POINTER(C.c_void_p)])
            _proto(lib, "gk_ipc_close", [_P])
            _proto(lib, "gk_ipc_free", [_P])
            _libs["cuda"] = lib
        return _libs["cuda"]


def cpu_lib():
    with _lock:
        if "cpu" not in _libs:
            lib = C.CDLL(build_cpu())
            ch, ds, sr = C.POINTER(GkChain), C.POINTER(GkDst), C.POINTER(GkSrc)
            _proto(lib, "gkcpu_version", [], C.c_char_p)
            _proto(lib, "gkcpu_compress", [ch, ds])
            _proto(lib, "gkcpu_dgc_compress", [ch, ds, C.c_uint64])
            _proto(lib, "gkcpu_scatter", [sr, _P, _P, C.c_float, C.c_int32])
            _proto(lib, "gkcpu_unscatter", [sr, _P, _P, C.c_int32])
            _proto(lib, "gkcpu_sgd", [_P, _P, _P, _P, C.c_int32, _P, _P])
            _proto(lib, "gkcpu_sgd_recv", [C.POINTER(GkRecvUpd), sr])
            _proto(lib, "gkcpu_lars", [_P, _P, _P, _P, C.c_int32, _P, C.c_int32, _P, _P, _P])
            _proto(lib, "gkcpu_grad_norm", [_P, _P, C.c_int32, C.c_float, _P])
            _proto(lib, "gkcpu_randomk_send", [ch, ds, C.c_uint64])
            _proto(lib, "gkcpu_bucket_means", [ch, _P])
            _proto(lib, "gkcpu_bucket_reconstruct", [ch, _P, _P])
            _libs["cpu"] = lib
        return _libs["cpu"]


def check(err: int, what: str = "gk") -> None:
    if err != 0:
        raise RuntimeError(f"{what}: native call failed with code {err}")
