"""soft-DTW CUDA path against the CPU oracle and the golden fixtures (GPU).

Tolerance (BASELINE.json): R and gradients within 1e-5 relative for float32,
measured against the reference's float64 numba-CPU recurrences (the oracle)."""
import glob
import os

import numpy as np
import pytest
import torch

import milnce_b200
from conftest import GOLDEN, rel_err
from milnce_b200 import SoftDTW
from milnce_b200.soft_dtw_cuda import sdtw_backward_raw, sdtw_forward_raw
from oracle import sdtw_oracle

pytestmark = pytest.mark.gpu
KINDS = milnce_b200._lib.DIST_KINDS


def _compare(x, y, gamma, bw, dist, grad_out=None, tol=1e-5, name=""):
    """Raw C-ABI forward+backward vs the float64 oracle: out, R, E, dX, dY."""
    xc, yc = x.cuda().contiguous(), y.cuda().contiguous()
    B = x.shape[0]
    go = torch.ones(B) if grad_out is None else grad_out
    out, R, aux = sdtw_forward_raw(xc, yc, gamma, 0.0 if bw is None else bw, KINDS[dist], want_R=True)
    dx, dy, E = sdtw_backward_raw(xc, yc, go.cuda().float().contiguous(), gamma,
                                  0.0 if bw is None else bw, KINDS[dist], aux, want_E=True)
    torch.cuda.synchronize()
    xn, yn = x.double().numpy(), y.double().numpy()
    g32 = float(np.float32(gamma))
    bwf = 0.0 if bw is None else float(bw)
    D = sdtw_oracle.dist_forward(xn, yn, dist)
    Rref = sdtw_oracle.dp_forward(D, g32, bwf)
    Eref = sdtw_oracle.dp_backward(D, Rref, g32, bwf)
    dxr, dyr = sdtw_oracle.dist_backward(xn, yn, D, go.double().numpy()[:, None, None] * Eref, dist)
    Rg = R.cpu().numpy().astype(np.float64)
    assert np.array_equal(np.isinf(Rg), np.isinf(Rref)), f"{name}: inf pattern of R differs"
    assert np.array_equal(np.sign(Rg[np.isinf(Rg)]), np.sign(Rref[np.isinf(Rref)]))
    fin = np.isfinite(Rref)
    e_R = np.max(np.abs(Rg[fin] - Rref[fin])) / max(np.max(np.abs(Rref[fin])), 1e-30)
    e_out = rel_err(out.cpu().numpy(), Rref[:, -2, -2])
    e_E = rel_err(E.cpu().numpy(), Eref)
    e_dx = rel_err(dx.cpu().numpy(), dxr)
    e_dy = rel_err(dy.cpu().numpy(), dyr)
    print(f"{name} {dist} g={gamma} bw={bw}: R {e_R:.2e} out {e_out:.2e} E {e_E:.2e} "
          f"dX {e_dx:.2e} dY {e_dy:.2e}")
    assert max(e_R, e_out) < tol, name
    # E and the input gradients see D through exp(./gamma): merely rounding D to float32 (which
    # the reference does too, its distance function runs in float32) perturbs E by about
    # ulp(|D|)/gamma.  For cosine |D| < e^2 and that stays below 1e-5; for the unnormalised
    # distances it does not, so the bound scales with |D|max there.
    gtol = max(tol, 8 * 2.0 ** -24 * float(np.max(np.abs(D))) / g32)
    assert max(e_E, e_dx, e_dy) < gtol, (name, gtol)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "sdtw_*.npz"))),
                         ids=os.path.basename)
def test_golden_fixture(path):
    g = np.load(path)
    x, y = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    bw = float(g["bandwidth"])
    go = torch.from_numpy(g["grad_out"])
    for dist in ("cosine", "negative_dot", "euclidean"):
        for gamma in (0.1, 1.0):
            _compare(x, y, gamma, bw if bw > 0 else None, dist, go, name=os.path.basename(path))
            # module API against what the executed reference returned (fp32-truncated R/E there)
            key = f"{dist}_g{gamma}_n0"
            xr = x.cuda().requires_grad_(True)
            yr = y.cuda().requires_grad_(True)
            mod = SoftDTW(True, gamma=gamma, bandwidth=(bw if bw > 0 else None), dist_func=dist)
            out = mod(xr, yr)
            (out * go.cuda()).sum().backward()
            scale = np.max(np.abs(g[key + "_out"])) + 1.0
            assert np.max(np.abs(out.detach().cpu().numpy() - g[key + "_out"])) < 5e-6 * scale
            gtol = max(2e-5, 8 * 2.0 ** -24 * scale / gamma)
            assert rel_err(xr.grad.cpu().numpy(), g[key + "_dx"]) < gtol
            assert rel_err(yr.grad.cpu().numpy(), g[key + "_dy"]) < gtol


@pytest.mark.parametrize("path", [p for p in sorted(glob.glob(os.path.join(GOLDEN, "sdtw_*.npz")))
                                  if "sq8" in p or "rand12" in p], ids=os.path.basename)
def test_normalize_matches_oracle(path):
    g = np.load(path)
    x, y = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    bw = float(g["bandwidth"])
    go = torch.from_numpy(g["grad_out"])
    for dist in ("cosine", "negative_dot"):
        for gamma in (0.1, 1.0):
            xr = x.cuda().requires_grad_(True)
            yr = y.cuda().requires_grad_(True)
            mod = SoftDTW(True, gamma=gamma, normalize=True, bandwidth=(bw if bw > 0 else None),
                          dist_func=dist)
            out = mod(xr, yr)
            (out * go.cuda()).sum().backward()
            o, dxr, dyr, _ = sdtw_oracle.softdtw(g["x"], g["y"], gamma=gamma, normalize=True,
                                                 bandwidth=(bw if bw > 0 else None), dist_func=dist,
                                                 grad_out=g["grad_out"])
            # the normalised value subtracts three O(|R|) terms: tolerance relative to those
            comp, _, _, _ = sdtw_oracle.softdtw(g["x"], g["x"], gamma=gamma, bandwidth=None
                                                if bw <= 0 else bw, dist_func=dist)
            scale = max(np.max(np.abs(comp)), np.max(np.abs(o)), 1.0)
            assert np.max(np.abs(out.detach().cpu().numpy() - o)) < 1e-5 * scale
            assert rel_err(xr.grad.cpu().numpy(), dxr) < 5e-5
            assert rel_err(yr.grad.cpu().numpy(), dyr) < 5e-5


@pytest.mark.parametrize("B,N,M,d,bw", [
    (16, 64, 64, 512, None),    # the BASELINE config's per-pair shape
    (4, 100, 37, 20, None),     # ragged, N > M
    (4, 37, 100, 20, 5),        # ragged, N < M, pruned: corner unreachable -> out = +inf
    (3, 65, 70, 33, 10),        # crosses tile and warp boundaries, d not a multiple of 4
    (2, 1, 1, 4, None),         # degenerate
    (2, 1, 9, 4, None),
    (2, 9, 1, 4, None),
])
@pytest.mark.parametrize("dist", ["cosine", "negative_dot", "euclidean"])
def test_random_shapes(B, N, M, d, bw, dist):
    g = torch.Generator().manual_seed(1234 + N * 3 + M)
    x = torch.randn(B, N, d, generator=g)
    y = torch.randn(B, M, d, generator=g)
    if dist == "euclidean":
        x, y = x * 0.2, y * 0.2   # exp(||x-y||) overflows float32 quickly otherwise
    if bw is not None and abs(N - M) > bw:
        # R[N, M] is +inf in the reference too; gradients are NaN there - compare R only
        xc, yc = x.cuda(), y.cuda()
        out, R, _ = sdtw_forward_raw(xc, yc, 0.1, bw, KINDS[dist], want_R=True)
        D = sdtw_oracle.dist_forward(x.double().numpy(), y.double().numpy(), dist)
        Rref = sdtw_oracle.dp_forward(D, float(np.float32(0.1)), float(bw))
        Rg = R.cpu().numpy().astype(np.float64)
        assert np.array_equal(np.isinf(Rg), np.isinf(Rref))
        fin = np.isfinite(Rref)
        assert np.max(np.abs(Rg[fin] - Rref[fin])) < 1e-5 * np.max(np.abs(Rref[fin]))
        assert bool(torch.isinf(out).all())
        return
    _compare(x, y, 0.1, bw, dist, name=f"B{B}N{N}M{M}d{d}")


def test_max_length_1024():
    g = torch.Generator().manual_seed(7)
    x = torch.randn(2, 1024, 8, generator=g)
    y = torch.randn(2, 1000, 8, generator=g)
    _compare(x, y, 1.0, None, "cosine", name="max-length")


def test_full_size_config():
    """BASELINE.json configs[2]: 256 pairs of 64x64, d=512, gamma=0.1, cosine, seed 1234
    (the reference's own profile() seed, soft_dtw_cuda.py:458)."""
    torch.manual_seed(1234)
    x = torch.randn(256, 64, 512)
    y = torch.randn(256, 64, 512)
    _compare(x, y, 0.1, None, "cosine", name="config3")
    # rand inputs as profile() uses (:427-428)
    _compare(torch.rand(64, 64, 512), torch.rand(64, 64, 512), 0.1, None, "cosine", name="config3-rand")


def test_linearity_in_grad_out():
    """d/dX is linear in grad_out: a size-independent property checked at full size."""
    torch.manual_seed(5)
    x = torch.randn(256, 64, 512, device="cuda")
    y = torch.randn(256, 64, 512, device="cuda")
    out, _, aux = sdtw_forward_raw(x, y, 0.1, 0.0, KINDS["cosine"])
    g1 = torch.rand(256, device="cuda")
    dx1, dy1, _ = sdtw_backward_raw(x, y, g1, 0.1, 0.0, KINDS["cosine"], aux)
    dx2, dy2, _ = sdtw_backward_raw(x, y, 3.0 * g1, 0.1, 0.0, KINDS["cosine"], aux)
    assert ((dx2 - 3.0 * dx1).norm() / dx2.norm()).item() < 1e-6
    assert ((dy2 - 3.0 * dy1).norm() / dy2.norm()).item() < 1e-6


def _forward_only(x, y, gamma, dist):
    out, R, _ = sdtw_forward_raw(x.cuda(), y.cuda(), gamma, 0.0, KINDS[dist], want_R=True)
    torch.cuda.synchronize()
    D = sdtw_oracle.dist_forward(x.double().numpy(), y.double().numpy(), dist)
    Rref = sdtw_oracle.dp_forward(D, float(np.float32(gamma)), 0.0)
    Rg = R.cpu().numpy().astype(np.float64)
    assert np.array_equal(np.isinf(Rg), np.isinf(Rref)), "inf pattern of R differs"
    fin = np.isfinite(Rref)
    e_R = np.max(np.abs(Rg[fin] - Rref[fin])) / np.max(np.abs(Rref[fin]))
    e_out = rel_err(out.cpu().numpy(), Rref[:, -2, -2])
    print(f"{dist} gamma={gamma}: R {e_R:.2e} out {e_out:.2e}")
    return e_R, e_out


@pytest.mark.parametrize("gamma", [1e-4, 1e-5, 1e-6])
def test_tiny_gamma_cosine_short_sequences(gamma):
    """N, M <= 64 with the gamma the reference's CDTW uses (1e-5, loss.py:74) and below: the per-pair
    kernel's scaled-pair exponents reach ~1e8 here and must neither meet the 'zero' sentinel nor
    leave the int range (gamma = 1e-6 is sent to the float64 general path by the host)."""
    g = torch.Generator().manual_seed(11)
    x = torch.randn(4, 64, 32, generator=g)
    y = torch.randn(4, 64, 32, generator=g)
    e_R, e_out = _forward_only(x, y, gamma, "cosine")
    assert max(e_R, e_out) < 1e-5


def test_negative_dot_beyond_the_short_kernel_range_falls_back_on_the_device():
    """|D| / gamma too large for the per-pair kernel's exponents (dots of tens at gamma = 1e-5): it
    raises its flag and the general kernels redo the batch in the same sdtw_forward call; a second
    call with benign values on the same aux must take the fast kernel again (flag cleared)."""
    g = torch.Generator().manual_seed(12)
    x = torch.randn(3, 48, 512, generator=g)
    y = torch.randn(3, 40, 512, generator=g)
    e_R, e_out = _forward_only(x, y, 1e-5, "negative_dot")
    assert max(e_R, e_out) < 1e-5
    _compare(x * 0.05, y * 0.05, 0.1, None, "negative_dot", name="after-fallback")


def test_unaligned_inputs_take_the_scalar_loaders():
    """A (B, N, d) view that starts 4 bytes into its storage is contiguous but not 16-byte aligned:
    it must be served (general path, scalar loads), not fault."""
    g = torch.Generator().manual_seed(13)
    B, N, M, d = 3, 24, 20, 16
    bx = torch.randn(B * N * d + 1, generator=g).cuda()
    by = torch.randn(B * M * d + 1, generator=g).cuda()
    x = bx[1:].view(B, N, d)
    y = by[1:].view(B, M, d)
    assert x.data_ptr() % 16 != 0 and x.is_contiguous()
    out, R, aux = sdtw_forward_raw(x, y, 0.5, 0.0, KINDS["cosine"], want_R=True)
    dx, dy, _ = sdtw_backward_raw(x, y, torch.ones(B, device="cuda"), 0.5, 0.0, KINDS["cosine"], aux)
    xa, ya = x.clone(), y.clone()
    out2, R2, aux2 = sdtw_forward_raw(xa, ya, 0.5, 0.0, KINDS["cosine"], want_R=True)
    dx2, dy2, _ = sdtw_backward_raw(xa, ya, torch.ones(B, device="cuda"), 0.5, 0.0, KINDS["cosine"], aux2)
    torch.cuda.synchronize()
    assert rel_err(out.cpu().numpy(), out2.cpu().numpy()) < 1e-5
    assert rel_err(dx.cpu().numpy(), dx2.cpu().numpy()) < 1e-5
    assert rel_err(dy.cpu().numpy(), dy2.cpu().numpy()) < 1e-5
