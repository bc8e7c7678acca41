"""ctypes binding of libb200head.so (the C ABI declared in include/b200head.h).

The shared library is built in-tree by :func:`build` (``nvcc`` for sm_100a) and
loaded lazily by :func:`load`.  There is no fallback: if the library cannot be
built or loaded, or a call fails, a ``RuntimeError`` is raised.
"""
from __future__ import annotations

import ctypes
import fcntl
import glob
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libb200head.so")
SOURCES = ("capi.cu", "milnce.cu", "sdtw.cu")
PUBLIC_HEADER = os.path.join(os.path.dirname(_HERE), "include", "b200head.h")

F32, BF16 = 0, 1
DIST_KINDS = {"cosine": 0, "negative_dot": 1, "euclidean": 2}

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC", "-t", "0",
]

_lib = None


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _deps():
    """Every file the library is compiled from: csrc/*.cu, csrc/*.cuh and the public header."""
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh"))) + [PUBLIC_HEADER]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in _deps())


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu into libb200head.so for sm_100a (cross-compiles without a GPU).

    Safe when several processes start at once (torchrun): the compile runs under an exclusive
    file lock into a temporary file that is renamed over the library, so nobody can load a
    half-written .so, and the ranks that waited for the lock find a fresh library and skip."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():   # another process built it while we waited
                return LIB_PATH
            if not os.path.exists(nvcc):
                raise RuntimeError("nvcc not found: cannot build libb200head.so")
            tmp = f"{LIB_PATH}.tmp.{os.getpid()}"
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp] + _sources()
            proc = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
            if proc.returncode != 0:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
            os.replace(tmp, LIB_PATH)
            if verbose:
                print(proc.stderr)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB_PATH


def _declare(lib):
    c = ctypes
    vp, fp, sz, i32, f32 = c.c_void_p, c.c_void_p, c.c_size_t, c.c_int, c.c_float
    lib.b200head_last_error.restype = c.c_char_p
    lib.b200head_last_error.argtypes = []
    lib.b200head_version.restype = i32
    lib.b200head_version.argtypes = []
    lib.b200head_launch_count.restype = c.c_ulonglong
    lib.b200head_launch_count.argtypes = []
    lib.milnce_workspace_bytes.restype = sz
    lib.milnce_workspace_bytes.argtypes = [i32, i32, i32, i32, i32]
    lib.milnce_forward.restype = i32
    lib.milnce_forward.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, fp, fp, fp, vp, sz, vp]
    lib.milnce_backward.restype = i32
    lib.milnce_backward.argtypes = [vp, vp, fp, fp, fp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32,
                                    vp, sz, vp]
    lib.milnce_forward_shard.restype = i32
    lib.milnce_forward_shard.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, fp, fp, fp, vp, sz, vp]
    lib.milnce_peer_slot_bytes.restype = sz
    lib.milnce_peer_slot_bytes.argtypes = [i32, i32, i32]
    lib.milnce_peer_publish.restype = i32
    lib.milnce_peer_publish.argtypes = [vp, vp, i32, i32, i32, vp, fp, vp]
    lib.milnce_forward_shard_peers.restype = i32
    lib.milnce_forward_shard_peers.argtypes = [c.POINTER(c.c_void_p), i32, i32, i32, i32, i32, i32, fp, fp, vp, sz, vp]
    lib.milnce_backward_shard_push.restype = i32
    lib.milnce_backward_shard_push.argtypes = [fp, fp, fp, c.POINTER(c.c_void_p), i32, i32, i32, i32, i32, i32, vp, sz, vp]
    lib.milnce_peer_reduce_dtext.restype = i32
    lib.milnce_peer_reduce_dtext.argtypes = [c.POINTER(c.c_void_p), i32, i32, i32, i32, i32, vp, i32, vp]
    lib.milnce_shard_combine_peers.restype = i32
    lib.milnce_shard_combine_peers.argtypes = [c.POINTER(c.c_void_p), i32, i32, fp, fp, fp, vp, vp]
    lib.milnce_shard_combine_scratch_bytes.restype = sz
    lib.milnce_shard_combine_scratch_bytes.argtypes = [i32, i32]
    lib.milnce_shard_combine.restype = i32
    lib.milnce_shard_combine.argtypes = [fp, i32, i32, fp, fp, fp, vp]
    lib.milnce_backward_shard.restype = i32
    lib.milnce_backward_shard.argtypes = [vp, vp, fp, fp, fp, vp, i32, vp, i32, i32, i32, i32, i32, i32, i32, i32,
                                          i32, vp, sz, vp]
    lib.sdtw_aux_bytes.restype = sz
    lib.sdtw_aux_bytes.argtypes = [i32, i32, i32, i32]
    lib.sdtw_forward.restype = i32
    lib.sdtw_forward.argtypes = [fp, fp, i32, i32, i32, i32, f32, f32, i32, fp, fp, vp, sz, vp]
    lib.sdtw_backward.restype = i32
    lib.sdtw_backward.argtypes = [fp, fp, fp, i32, i32, i32, i32, f32, f32, i32, vp, sz, fp, fp, fp,
                                  vp]
    lib.sdtw_forward_pairs.restype = i32
    lib.sdtw_forward_pairs.argtypes = [fp, i32, fp, i32, vp, vp, i32, i32, i32, i32, f32, f32, i32, fp, fp, vp, sz, vp]
    lib.sdtw_backward_pairs.restype = i32
    lib.sdtw_backward_pairs.argtypes = [fp, i32, fp, i32, vp, vp, i32, fp, i32, i32, i32, f32, f32, i32, vp, sz, fp,
                                        fp, fp, vp]
    lib.sdtw_forward_noalloc.restype = i32
    lib.sdtw_forward_noalloc.argtypes = [fp, fp, i32, i32, i32, i32, f32, f32, i32, fp, fp, vp]
    lib.b200head_gemm_tn_f16.restype = i32
    lib.b200head_gemm_tn_f16.argtypes = [vp, vp, fp, i32, i32, i32, i32, i32, i32, f32, vp]
    lib.b200head_gemm_at_f16.restype = i32
    lib.b200head_gemm_at_f16.argtypes = [vp, vp, fp, i32, i32, i32, i32, i32, i32, f32, vp]
    lib.b200head_gemm2_f16.restype = i32
    lib.b200head_gemm2_f16.argtypes = [vp, vp, fp, i32, i32, i32, i32, i32, i32, f32, i32, vp]


def load():
    """Return the loaded library, building it first if it is missing or stale."""
    global _lib
    if _lib is None:
        alt = os.environ.get("B200HEAD_LIB")  # developer hook: an experimental build of the same sources
        if not alt:
            build()
        lib = ctypes.CDLL(alt or LIB_PATH)
        _declare(lib)
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200head_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def stream_ptr(device=None):
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
