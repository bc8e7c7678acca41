"""CPU oracle for SoftDTW.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module; the product
path never does.

Restates, in float64:

* the three distance functions of this fork (reference
  ``soft_dtw_cuda.py:326-363``): cosine = exp(1 - cos_sim), euclidean =
  exp(||x - y||), negative_dot = -<x, y>, and their input gradients (the
  reference gets those from torch autograd; here they are written out);
* the soft-DTW recurrences (``soft_dtw_cuda.py:185-240``) through the C
  restatement in ``oracle/sdtw_dp.c`` (or pure-Python loops for tiny cases);
* the ``normalize=True`` combination (``soft_dtw_cuda.py:376-383``).

Pinned against the executed reference (numba-CPU path, the f64-internal one,
SURVEY.md F4/F5) by ``oracle/make_golden.py`` -> ``tests/golden/sdtw_*.npz``.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libsdtw_oracle.so")
_lib = None

DIST_FUNCS = ("cosine", "negative_dot", "euclidean")
COS_EPS = 1e-8  # torch.nn.functional.cosine_similarity default (soft_dtw_cuda.py:347)


def build(force: bool = False) -> str:
    """Compile oracle/sdtw_dp.c with gcc (idempotent)."""
    src = os.path.join(_HERE, "sdtw_dp.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(
            ["gcc", "-O2", "-fPIC", "-shared", "-o", _LIB_PATH, src, "-lm"], cwd=_HERE
        )
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        dp = ctypes.POINTER(ctypes.c_double)
        lib.sdtw_oracle_forward.argtypes = [dp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                            ctypes.c_double, ctypes.c_double, dp]
        lib.sdtw_oracle_forward.restype = None
        lib.sdtw_oracle_backward.argtypes = [dp, dp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_double, ctypes.c_double, dp]
        lib.sdtw_oracle_backward.restype = None
        _lib = lib
    return _lib


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


# ----------------------------------------------------------------------------
# dynamic programme
# ----------------------------------------------------------------------------
def dp_forward(D: np.ndarray, gamma: float, bandwidth: float = 0.0) -> np.ndarray:
    """R (B,N+2,M+2) float64 - soft_dtw_cuda.py:185-206."""
    D = np.ascontiguousarray(D, dtype=np.float64)
    b, n, m = D.shape
    R = np.empty((b, n + 2, m + 2), dtype=np.float64)
    _load().sdtw_oracle_forward(_ptr(D), b, n, m, float(gamma), float(bandwidth), _ptr(R))
    return R


def dp_backward(D: np.ndarray, R: np.ndarray, gamma: float, bandwidth: float = 0.0) -> np.ndarray:
    """E[1:N+1,1:M+1] = d out / d D, float64 - soft_dtw_cuda.py:209-240.  R is copied."""
    D = np.ascontiguousarray(D, dtype=np.float64)
    R = np.array(R, dtype=np.float64, order="C", copy=True)
    b, n, m = D.shape
    E = np.empty((b, n + 2, m + 2), dtype=np.float64)
    _load().sdtw_oracle_backward(_ptr(D), _ptr(R), b, n, m, float(gamma), float(bandwidth), _ptr(E))
    return E[:, 1:n + 1, 1:m + 1].copy()


def dp_forward_py(D: np.ndarray, gamma: float, bandwidth: float = 0.0) -> np.ndarray:
    """Pure-Python loops, tiny cases only (cross-check of the C code)."""
    D = np.asarray(D, dtype=np.float64)
    b, n, m = D.shape
    R = np.full((b, n + 2, m + 2), np.inf)
    R[:, 0, 0] = 0.0
    for bb in range(b):
        for j in range(1, m + 1):
            for i in range(1, n + 1):
                if 0 < bandwidth < abs(i - j):
                    continue
                r = -np.array([R[bb, i - 1, j - 1], R[bb, i - 1, j], R[bb, i, j - 1]]) / gamma
                rmax = r.max()
                R[bb, i, j] = D[bb, i - 1, j - 1] - gamma * (np.log(np.exp(r - rmax).sum()) + rmax)
    return R


# ----------------------------------------------------------------------------
# distance functions and their gradients
# ----------------------------------------------------------------------------
def dist_forward(x: np.ndarray, y: np.ndarray, dist_func: str) -> np.ndarray:
    """D (B,N,M) float64 - soft_dtw_cuda.py:326-363."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    dot = np.einsum("bnd,bmd->bnm", x, y)
    if dist_func == "negative_dot":
        return -dot
    if dist_func == "cosine":
        nx = np.maximum(np.linalg.norm(x, axis=2), COS_EPS)
        ny = np.maximum(np.linalg.norm(y, axis=2), COS_EPS)
        return np.exp(1.0 - dot / (nx[:, :, None] * ny[:, None, :]))
    if dist_func == "euclidean":
        diff = x[:, :, None, :] - y[:, None, :, :]
        return np.exp(np.sqrt((diff * diff).sum(axis=3)))
    raise AttributeError(f"unknown dist_func {dist_func!r}")


def dist_backward(x, y, D, gD, dist_func: str):
    """(dX, dY) given gD = d loss / d D, float64 (autograd of :326-363 written out)."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    if dist_func == "negative_dot":
        return -np.einsum("bnm,bmd->bnd", gD, y), -np.einsum("bnm,bnd->bmd", gD, x)
    if dist_func == "cosine":
        rx = np.linalg.norm(x, axis=2)
        ry = np.linalg.norm(y, axis=2)
        nx = np.maximum(rx, COS_EPS)
        ny = np.maximum(ry, COS_EPS)
        xh = x / nx[:, :, None]
        yh = y / ny[:, :, None]
        c = np.einsum("bnd,bmd->bnm", xh, yh)
        gc = -gD * D  # D = exp(1 - c)
        # torch clamps the norms under no_grad (ATen Distance.cpp), so the norm's
        # own derivative is that of the true norm: d n / d x = x / ||x|| (0 at x = 0).
        with np.errstate(divide="ignore", invalid="ignore"):
            ux = np.where(rx[:, :, None] > 0, x / rx[:, :, None], 0.0)
            uy = np.where(ry[:, :, None] > 0, y / ry[:, :, None], 0.0)
        dx = (np.einsum("bnm,bmd->bnd", gc, yh)
              - (gc * c).sum(axis=2)[:, :, None] * ux) / nx[:, :, None]
        dy = (np.einsum("bnm,bnd->bmd", gc, xh)
              - (gc * c).sum(axis=1)[:, :, None] * uy) / ny[:, :, None]
        return dx, dy
    if dist_func == "euclidean":
        diff = x[:, :, None, :] - y[:, None, :, :]
        r = np.sqrt((diff * diff).sum(axis=3))
        with np.errstate(divide="ignore", invalid="ignore"):
            w = gD * D / r  # NaN at r == 0, like torch's sqrt backward
        dx = np.einsum("bnm,bnmd->bnd", w, diff)
        dy = -np.einsum("bnm,bnmd->bmd", w, diff)
        return dx, dy
    raise AttributeError(f"unknown dist_func {dist_func!r}")


# ----------------------------------------------------------------------------
# module-level restatement
# ----------------------------------------------------------------------------
def softdtw(x, y, gamma=1.0, normalize=False, bandwidth=None, dist_func="cosine", grad_out=None):
    """float64 out (B,), dX, dY, R (B,N+2,M+2) for SoftDTW(...).forward(x, y).

    Follows soft_dtw_cuda.py:365-386.  gamma is rounded through float32 because
    the reference's CPU path casts it to the input dtype (soft_dtw_cuda.py:252).
    R is returned for the plain (non-normalised) case only.
    """
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    if x.shape[0] != y.shape[0] or x.shape[2] != y.shape[2]:
        raise ValueError("batch and feature dims must match")
    g32 = float(np.float32(gamma))
    bw = 0.0 if bandwidth is None else float(np.float32(bandwidth))
    b = x.shape[0]
    go = np.ones(b) if grad_out is None else np.asarray(grad_out, dtype=np.float64)

    def one(xa, ya, g):
        D = dist_forward(xa, ya, dist_func)
        R = dp_forward(D, g32, bw)
        out = R[:, -2, -2].copy()
        E = dp_backward(D, R, g32, bw)
        dxa, dya = dist_backward(xa, ya, D, g[:, None, None] * E, dist_func)
        return out, dxa, dya, R

    if not normalize:
        return one(x, y, go)
    if x.shape[1] != y.shape[1]:
        raise ValueError("normalize=True needs equal sequence lengths (torch.cat at :378-379)")
    out_xy, dx1, dy1, _ = one(x, y, go)
    out_xx, dxa, dxb, _ = one(x, x, -0.5 * go)
    out_yy, dya, dyb, _ = one(y, y, -0.5 * go)
    return out_xy - 0.5 * (out_xx + out_yy), dx1 + dxa + dxb, dy1 + dya + dyb, None
