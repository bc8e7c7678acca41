"""sdtw_forward_noalloc: the forward for callers that cannot allocate device memory.

Some users of this library only ever hold input and output pointers: a binding from
another runtime with no device allocator of its own, a process whose memory pool
already owns the whole GPU, or a loop that must not pay the device-wide stall of
cudaMalloc/cudaFree.  For them the library carries its own 8 MiB scratch buffer.

The entry point is driven through the C ABI (ctypes) and compared with the regular
sdtw_forward entry point (caller-allocated aux) on consecutive calls.  Later calls are
smaller than the first, so they run on top of the first call's leftovers in the shared
buffer - the results must not depend on them.  Tolerance: 1e-5 relative, the same as the
other soft-DTW tests.  What the comparison means depends on the shape:
  * 256 x (64 x 64) x 512, cosine: the regular entry point takes the fused per-pair
    kernel there while the no-alloc one always runs the distance + wavefront kernels of
    the general path, so this is a cross-check between two kernel families (1e-5);
  * shapes the fused kernel does not cover (N = 80, or euclidean): both entry points run
    the very same kernels, and the outputs must be bit-identical.
The first call is also checked against the CPU oracle directly.  This file is named to be
collected last.
"""
import ctypes

import numpy as np
import pytest
import torch

import milnce_b200
from conftest import rel_err

pytestmark = pytest.mark.gpu
KINDS = milnce_b200._lib.DIST_KINDS
MAX_BYTES = 8 << 20  # B200HEAD_SDTW_NOALLOC_MAX_BYTES in include/b200head.h


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _noalloc(x, y, gamma, bw, kind):
    lib = milnce_b200._lib.load()
    B, N, d = x.shape
    M = y.shape[1]
    out = torch.full((B,), float("nan"), device="cuda")
    R = torch.full((B, N + 2, M + 2), float("nan"), device="cuda")
    rc = lib.sdtw_forward_noalloc(_ptr(x), _ptr(y), B, N, M, d, ctypes.c_float(gamma),
                                  ctypes.c_float(bw), kind, _ptr(R), _ptr(out),
                                  milnce_b200._lib.stream_ptr())
    milnce_b200._lib.check(rc, "sdtw_forward_noalloc")
    torch.cuda.synchronize()
    return out, R


def _regular(x, y, gamma, bw, kind):
    from milnce_b200.soft_dtw_cuda import sdtw_forward_raw
    out, R, _ = sdtw_forward_raw(x, y, gamma, bw, kind, want_R=True)
    torch.cuda.synchronize()
    return out, R


def _same(a, b):
    a, b = a.cpu().numpy().astype(np.float64), b.cpu().numpy().astype(np.float64)
    assert np.array_equal(np.isinf(a), np.isinf(b))
    fin = np.isfinite(b)
    return rel_err(a[fin], b[fin])


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    g = torch.Generator().manual_seed(2024)
    # call 1: the BASELINE soft-DTW shape (4.3 MiB of the 8 MiB buffer)
    x1 = torch.randn(256, 64, 512, generator=g).cuda()
    y1 = torch.randn(256, 64, 512, generator=g).cuda()
    # call 2: smaller, ragged and pruned, straight after - reuses the dirty buffer
    x2 = torch.randn(5, 37, 24, generator=g).cuda()
    y2 = torch.randn(5, 41, 24, generator=g).cuda()
    # call 3: longer than the fused kernels' 64 frames - both entry points run the same kernels
    x3 = torch.randn(6, 80, 32, generator=g).cuda()
    y3 = torch.randn(6, 72, 32, generator=g).cuda()
    calls = [(x1, y1, 0.1, 0.0, KINDS["cosine"]), (x2, y2, 1.0, 6.0, KINDS["negative_dot"]),
             (x3, y3, 0.5, 0.0, KINDS["cosine"])]
    got = [_noalloc(*c) for c in calls]        # calls in a row, nothing in between
    for c, (out, R) in zip(calls, got):
        ref_out, ref_R = _regular(*c)
        assert _same(out, ref_out) < 1e-5
        assert _same(R, ref_R) < 1e-5
    # same kernels on both sides: bit-identical
    out3, R3 = got[2]
    ref_out3, ref_R3 = _regular(*calls[2])
    assert torch.equal(out3, ref_out3) and torch.equal(R3, ref_R3)
    # and the first call against the CPU oracle itself (float64 recurrence, soft_dtw_cuda.py:185-206)
    from oracle import sdtw_oracle
    sub = slice(0, 16)
    D = sdtw_oracle.dist_forward(x1[sub].double().cpu().numpy(), y1[sub].double().cpu().numpy(), "cosine")
    Rref = sdtw_oracle.dp_forward(D, float(np.float32(0.1)), 0.0)
    assert rel_err(got[0][0][sub].cpu().numpy(), Rref[:, -2, -2]) < 1e-5


def test_noalloc_euclidean_is_bit_identical_to_regular():
    """euclidean never takes the fused kernel: the two entry points run the same code."""
    g = torch.Generator().manual_seed(7)
    x = (torch.randn(9, 33, 20, generator=g) * 0.3).cuda()
    y = (torch.randn(9, 29, 20, generator=g) * 0.3).cuda()
    out, R = _noalloc(x, y, 0.7, 0.0, KINDS["euclidean"])
    ref_out, ref_R = _regular(x, y, 0.7, 0.0, KINDS["euclidean"])
    assert torch.equal(out, ref_out) and torch.equal(R, ref_R)


def test_noalloc_rejects_sizes_beyond_the_builtin_buffer():
    lib = milnce_b200._lib.load()
    B, N, M, d = 1024, 64, 64, 8          # D alone is 16 MiB > 8 MiB
    assert 4 * B * N * M > MAX_BYTES
    x = torch.zeros(B, N, d, device="cuda")
    y = torch.zeros(B, M, d, device="cuda")
    out = torch.zeros(B, device="cuda")
    rc = lib.sdtw_forward_noalloc(_ptr(x), _ptr(y), B, N, M, d, ctypes.c_float(1.0),
                                  ctypes.c_float(0.0), 0, _ptr(None), _ptr(out),
                                  milnce_b200._lib.stream_ptr())
    assert rc == -4
    msg = lib.b200head_last_error().decode()
    assert "needs" in msg and "sdtw_forward" in msg


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_noalloc_uses_the_scratch_of_the_current_device():
    """The built-in buffer exists once per GPU and its address is looked up per device: a call on the second GPU
    after one on the first must use the second GPU's own copy (a cached pointer of device 0 would fault there or
    give wrong numbers)."""
    lib = milnce_b200._lib.load()
    g = torch.Generator().manual_seed(99)
    xs = torch.randn(7, 48, 16, generator=g)
    ys = torch.randn(7, 50, 16, generator=g)
    outs = []
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            x, y = xs.cuda(dev), ys.cuda(dev)
            out = torch.full((7,), float("nan"), device=f"cuda:{dev}")
            rc = lib.sdtw_forward_noalloc(_ptr(x), _ptr(y), 7, 48, 50, 16, ctypes.c_float(0.5), ctypes.c_float(0.0),
                                          KINDS["euclidean"], _ptr(None), _ptr(out),
                                          milnce_b200._lib.stream_ptr(torch.device("cuda", dev)))
            milnce_b200._lib.check(rc, "sdtw_forward_noalloc")
            torch.cuda.synchronize(dev)
            outs.append(out.cpu())
    assert torch.isfinite(outs[0]).all()
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
