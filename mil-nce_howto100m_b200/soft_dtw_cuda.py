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


def _iptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class _SoftDTWPairsFunction(torch.autograd.Function):
    """soft-DTW of index pairs (sdtw_forward_pairs / sdtw_backward_pairs of the C ABI): pair p compares
    X[ix[p]] with Y[iy[p]].  The gradient of a sequence is the sum over the pairs it takes part in."""

    @staticmethod
    def forward(ctx, X, Y, ix, iy, gamma, bandwidth, dist_kind):
        lib = _lib.load()
        x = X.detach().contiguous()
        y = Y.detach().contiguous()
        Bx, N, d = x.shape
        By, M, _ = y.shape
        ix = None if ix is None else ix.to(device=x.device, dtype=torch.int32).contiguous()
        iy = None if iy is None else iy.to(device=x.device, dtype=torch.int32).contiguous()
        P = ix.numel() if ix is not None else (iy.numel() if iy is not None else Bx)
        dev = x.device
        aux_bytes = lib.sdtw_aux_bytes(P, N, M, d)
        aux = torch.empty(aux_bytes, dtype=torch.uint8, device=dev)
        out = torch.empty(P, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.sdtw_forward_pairs(_ptr(x), Bx, _ptr(y), By, _iptr(ix), _iptr(iy), P, N, M, d, float(gamma),
                                        float(bandwidth), int(dist_kind), _ptr(None), _ptr(out), _ptr(aux), aux_bytes,
                                        _lib.stream_ptr(dev))
        _lib.check(rc, "sdtw_forward_pairs")
        ctx.save_for_backward(x, y)
        ctx.idx = (ix, iy, P)
        ctx.aux = aux
        ctx.cfg = (float(gamma), float(bandwidth), int(dist_kind))
        return out

    @staticmethod
    def backward(ctx, grad_output):
        lib = _lib.load()
        x, y = ctx.saved_tensors
        ix, iy, P = ctx.idx
        gamma, bandwidth, dist_kind = ctx.cfg
        Bx, N, d = x.shape
        By, M, _ = y.shape
        g = grad_output.detach().to(torch.float32).contiguous()
        dx = torch.empty_like(x)
        dy = torch.empty_like(y)
        dev = x.device
        with torch.cuda.device(dev):
            rc = lib.sdtw_backward_pairs(_ptr(x), Bx, _ptr(y), By, _iptr(ix), _iptr(iy), P, _ptr(g), N, M, d, gamma,
                                         bandwidth, dist_kind, _ptr(ctx.aux), ctx.aux.numel(), _ptr(None), _ptr(dx),
                                         _ptr(dy), _lib.stream_ptr(dev))
        _lib.check(rc, "sdtw_backward_pairs")
        return dx, dy, None, None, None, None, None


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
            # The reference stacks cat([X, X, Y]) against cat([Y, X, Y]) (soft_dtw_cuda.py:378-379): three
            # batches of pairs.  Here each batch reads its operands where they are - no stacked copies.
            out_xy = _SoftDTWFunction.apply(X, Y, self.gamma, self.bandwidth, kind)
            out_xx = _SoftDTWFunction.apply(X, X, self.gamma, self.bandwidth, kind)
            out_yy = _SoftDTWFunction.apply(Y, Y, self.gamma, self.bandwidth, kind)
            return out_xy - 1 / 2 * (out_xx + out_yy)
        return _SoftDTWFunction.apply(X, Y, self.gamma, self.bandwidth, kind)

    def forward_pairs(self, X, Y, ix=None, iy=None):
        """soft-DTW of the pairs (X[ix[p]], Y[iy[p]]) -> (P,).  Not in the reference API: it replaces the
        expand(...).reshape(...) / repeat(...) operand copies its composite losses make before calling
        forward (loss.py:29, 101-105); normalize is not applied."""
        self._check_pairs(X, Y)
        kind = _lib.DIST_KINDS[self.dist_func]
        return _SoftDTWPairsFunction.apply(X, Y, ix, iy, self.gamma, self.bandwidth, kind)

    def _check_pairs(self, x, y):
        if x.dim() != 3 or y.dim() != 3:
            raise ValueError("X and Y must be (Bx, N, d) and (By, M, d)")
        if x.shape[2] != y.shape[2]:
            raise ValueError("feature dimensions differ")
        if not self.use_cuda:
            raise RuntimeError("SoftDTW(use_cuda=False): this package has no CPU path")
        if x.shape[1] > 1024 or y.shape[1] > 1024:
            raise RuntimeError("SoftDTW: sequence length > 1024 is not supported")
        if not (x.is_cuda and y.is_cuda):
            raise RuntimeError("SoftDTW runs on CUDA tensors only")
        if x.dtype != torch.float32 or y.dtype != torch.float32:
            raise TypeError("SoftDTW supports float32 inputs")
