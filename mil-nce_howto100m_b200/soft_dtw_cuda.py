"""Drop-in for the reference's ``soft_dtw_cuda.py::SoftDTW`` (soft_dtw_cuda.py:274-386).

``SoftDTW(use_cuda, gamma, normalize, bandwidth, dist_func)(X, Y)`` -> (B,), with
gradients w.r.t. X and Y.  The distance function, the wavefront dynamic programme
and both backward passes run in libb200head.so.

Deliberate, documented deviations from the reference:
* ``use_cuda=False`` and sequence lengths > 1024 raise instead of running the
  numba-CPU path (soft_dtw_cuda.py:318-323): this package has no CPU path;
* float32 only (the reference's CUDA path allocates ``torch.cuda.FloatTensor``,
  soft_dtw_cuda.py:125);
* R is carried in float64 inside the kernel, so gradients agree with the
  reference's float64 numba-CPU path rather than with its float32-storing
  numba-CUDA kernels.
Mirrored quirks: ``dist_func=None`` leaves ``self.dist_func`` unset, so ``forward``
raises ``AttributeError`` like the reference (soft_dtw_cuda.py:297-304);
``'negative_cosine'`` raises ``AttributeError`` at construction.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def sdtw_forward_raw(x, y, gamma, bandwidth, dist_kind, want_R=False):
    """One sdtw_forward call through the C ABI.  Returns (out, R or None, aux)."""
    lib = _lib.load()
    B, N, d = x.shape
    M = y.shape[1]
    dev = x.device
    aux_bytes = lib.sdtw_aux_bytes(B, N, M, d)
    aux = torch.empty(aux_bytes, dtype=torch.uint8, device=dev)
    out = torch.empty(B, dtype=torch.float32, device=dev)
    R = torch.empty(B, N + 2, M + 2, dtype=torch.float32, device=dev) if want_R else None
    with torch.cuda.device(dev):
        rc = lib.sdtw_forward(_ptr(x), _ptr(y), B, N, M, d, float(gamma), float(bandwidth),
                              int(dist_kind), _ptr(R), _ptr(out), _ptr(aux), aux_bytes,
                              _lib.stream_ptr(dev))
    _lib.check(rc, "sdtw_forward")
    return out, R, aux


def sdtw_backward_raw(x, y, grad_out, gamma, bandwidth, dist_kind, aux, want_E=False):
    """One sdtw_backward call through the C ABI.  Returns (dx, dy, E or None)."""
    lib = _lib.load()
    B, N, d = x.shape
    M = y.shape[1]
    dev = x.device
    dx = torch.empty_like(x)
    dy = torch.empty_like(y)
    E = torch.empty(B, N, M, dtype=torch.float32, device=dev) if want_E else None
    with torch.cuda.device(dev):
        rc = lib.sdtw_backward(_ptr(x), _ptr(y), _ptr(grad_out), B, N, M, d, float(gamma),
                               float(bandwidth), int(dist_kind), _ptr(aux), aux.numel(), _ptr(E),
                               _ptr(dx), _ptr(dy), _lib.stream_ptr(dev))
    _lib.check(rc, "sdtw_backward")
    return dx, dy, E


class _SoftDTWFunction(torch.autograd.Function):
    """Fused replacement for dist_func + _SoftDTWCUDA (soft_dtw_cuda.py:115-175, 326-363)."""

    @staticmethod
    def forward(ctx, X, Y, gamma, bandwidth, dist_kind):
        x = X.detach().contiguous()
        y = Y.detach().contiguous()
        out, _, aux = sdtw_forward_raw(x, y, gamma, bandwidth, dist_kind)
        ctx.save_for_backward(x, y)
        ctx.aux = aux
        ctx.cfg = (float(gamma), float(bandwidth), int(dist_kind))
        return out

    @staticmethod
    def backward(ctx, grad_output):
        x, y = ctx.saved_tensors
        gamma, bandwidth, dist_kind = ctx.cfg
        g = grad_output.detach().to(torch.float32).contiguous()
        dx, dy, _ = sdtw_backward_raw(x, y, g, gamma, bandwidth, dist_kind, ctx.aux)
        return dx, dy, None, None, None


class SoftDTW(torch.nn.Module):
    """The soft-DTW module of the reference, backed by sm_100a kernels."""

    def __init__(self, use_cuda, gamma=1.0, normalize=False, bandwidth=None, dist_func=None):
        super().__init__()
        self.normalize = normalize
        self.gamma = gamma
        self.bandwidth = 0 if bandwidth is None else float(bandwidth)
        self.use_cuda = use_cuda
        if dist_func in _lib.DIST_KINDS:
            self.dist_func = dist_func
            self.dist_kind = _lib.DIST_KINDS[dist_func]
        elif dist_func == "negative_cosine":
            # the reference looks up a method that does not exist (soft_dtw_cuda.py:300)
            raise AttributeError("type object 'SoftDTW' has no attribute '_negative_cosine_dist_func'")
        # any other value (including None): dist_func stays unset, forward raises AttributeError

    def _check(self, x, y):
        if x.dim() != 3 or y.dim() != 3:
            raise ValueError("X and Y must be (B, N, d) and (B, M, d)")
        bx, lx, dx = x.shape
        by, ly, dy = y.shape
        if bx != by:
            raise ValueError("batch sizes differ")      # reference: assert, :313
        if dx != dy:
            raise ValueError("feature dimensions differ")  # reference: assert, :314
        if not self.use_cuda:
            raise RuntimeError("SoftDTW(use_cuda=False): this package has no CPU path")
        if lx > 1024 or ly > 1024:
            raise RuntimeError("SoftDTW: sequence length > 1024 is not supported (the reference "
                               "falls back to CPU here; this package has no CPU path)")
        if not (x.is_cuda and y.is_cuda):
            raise RuntimeError("SoftDTW runs on CUDA tensors only")
        if x.dtype != torch.float32 or y.dtype != torch.float32:
            raise TypeError("SoftDTW supports float32 inputs")

    def forward(self, X, Y):
        self._check(X, Y)
        kind = _lib.DIST_KINDS[self.dist_func]  # AttributeError when dist_func was None
        if self.normalize:
            if X.shape[1] != Y.shape[1]:
                raise RuntimeError("normalize=True needs equal sequence lengths (torch.cat, :378)")
            x = torch.cat([X, X, Y])
            y = torch.cat([Y, X, Y])
            out = _SoftDTWFunction.apply(x, y, self.gamma, self.bandwidth, kind)
            out_xy, out_xx, out_yy = torch.split(out, X.shape[0])
            return out_xy - 1 / 2 * (out_xx + out_yy)
        return _SoftDTWFunction.apply(X, Y, self.gamma, self.bandwidth, kind)
