"""Pair-list soft-DTW (sdtw_forward_pairs / sdtw_backward_pairs) and the composite soft-DTW losses of
reference loss.py:20-134 against fixtures minted from the executed reference (oracle/make_golden.py:
CDTW, SDTW_3, SDTW_CIDM run with their SoftDTW on the numba-CPU path, float64)."""
import os

import numpy as np
import pytest
import torch

import milnce_b200
from conftest import GOLDEN, rel_err
from milnce_b200 import SoftDTW

pytestmark = pytest.mark.gpu


class Args:
    rank = 1


@pytest.mark.parametrize("dist", ["cosine", "negative_dot", "euclidean"])
@pytest.mark.parametrize("Bx,By,N,M,d,P", [(5, 7, 12, 9, 16, 23), (4, 4, 64, 64, 64, 16), (3, 6, 70, 40, 20, 11)])
def test_pairs_equal_explicit_gather(Bx, By, N, M, d, P, dist):
    """forward_pairs(X, Y, ix, iy) == forward(X[ix], Y[iy]); the gradient of a sequence is the sum of the
    gradients of the pairs it takes part in (what autograd gives through the gather)."""
    g = torch.Generator().manual_seed(Bx * 100 + By * 10 + P)
    scale = 0.2 if dist == "euclidean" else 1.0
    X = (torch.randn(Bx, N, d, generator=g) * scale).cuda().requires_grad_(True)
    Y = (torch.randn(By, M, d, generator=g) * scale).cuda().requires_grad_(True)
    ix = torch.randint(0, Bx, (P,), generator=g).cuda()
    iy = torch.randint(0, By, (P,), generator=g).cuda()
    go = (torch.rand(P, generator=g) + 0.5).cuda()
    mod = SoftDTW(True, gamma=0.5, dist_func=dist)
    out = mod.forward_pairs(X, Y, ix, iy)
    (out * go).sum().backward()
    dX, dY = X.grad.clone(), Y.grad.clone()
    X.grad = Y.grad = None
    ref = mod(X[ix], Y[iy])
    (ref * go).sum().backward()
    torch.cuda.synchronize()
    assert rel_err(out.detach().cpu().numpy(), ref.detach().cpu().numpy()) < 1e-6
    assert rel_err(dX.cpu().numpy(), X.grad.cpu().numpy()) < 1e-5
    assert rel_err(dY.cpu().numpy(), Y.grad.cpu().numpy()) < 1e-5


def test_pairs_identity_lists_and_errors():
    g = torch.Generator().manual_seed(3)
    X = torch.randn(6, 10, 8, generator=g).cuda()
    Y = torch.randn(6, 10, 8, generator=g).cuda()
    mod = SoftDTW(True, gamma=1.0, dist_func="cosine")
    a = mod.forward_pairs(X, Y)                       # no lists: the plain batched call
    b = mod(X, Y)
    assert torch.equal(a, b)
    c = mod.forward_pairs(X, Y, torch.arange(6).cuda(), None)
    assert torch.allclose(c, b, rtol=1e-6, atol=0)
    with pytest.raises(RuntimeError, match="exactly P"):
        mod.forward_pairs(X, Y, torch.zeros(4, dtype=torch.int32).cuda(), None)   # P = 4 but Y holds 6


def _fix():
    return np.load(os.path.join(GOLDEN, "composite_losses.npz"))


def test_cdtw_matches_executed_reference():
    f = _fix()
    v = torch.from_numpy(f["cdtw_video"]).cuda().requires_grad_(True)
    t = torch.from_numpy(f["cdtw_text"]).cuda().requires_grad_(True)
    out = milnce_b200.CDTW(Args())(v, t)
    out.sum().backward()
    torch.cuda.synchronize()
    assert out.shape == f["cdtw_out"].shape
    ref_out = float(f["cdtw_out"].reshape(-1)[0])
    assert abs(out.item() - ref_out) < 1e-5 * max(1.0, abs(ref_out))
    # gamma = 1e-5: the float32 rounding of the distances alone perturbs the soft-min weights by
    # ulp(D) / gamma ~ 2e-2 (the bound the other soft-DTW tests use: 8 * 2^-24 * max|D| / gamma, D <= e^2)
    gtol = 8 * 2.0 ** -24 * float(np.exp(2.0)) / 1e-5
    e_v, e_t = rel_err(v.grad.cpu().numpy(), f["cdtw_dvideo"]), rel_err(t.grad.cpu().numpy(), f["cdtw_dtext"])
    print(f"CDTW gradient errors {e_v:.2e} {e_t:.2e} (bound {gtol:.2e})")
    assert e_v < gtol and e_t < gtol


def test_sdtw3_matches_executed_reference():
    f = _fix()
    v = torch.from_numpy(f["s3_video"]).cuda().requires_grad_(True)
    t = torch.from_numpy(f["s3_text"]).cuda().requires_grad_(True)
    l1, l2, l3 = milnce_b200.SDTW_3(Args())(v, t)
    (l1 + 2.0 * l2 + 3.0 * l3).backward()
    torch.cuda.synchronize()
    got = np.array([l1.item(), l2.item(), l3.item()])
    # losses 1 and 3 are logsumexp(neg) - pos with pos the dominant term: near zero, compare absolutely
    assert np.max(np.abs(got - f["s3_losses"])) < 2e-5
    assert rel_err(v.grad.cpu().numpy(), f["s3_dvideo"]) < 1e-4
    assert rel_err(t.grad.cpu().numpy(), f["s3_dtext"]) < 1e-4


def test_sdtw_cidm_matches_executed_reference():
    f = _fix()
    v = torch.from_numpy(f["cidm_video"]).cuda().requires_grad_(True)
    t = torch.from_numpy(f["cidm_text"]).cuda().requires_grad_(True)
    start = torch.from_numpy(f["cidm_start"]).float().cuda()
    end = torch.from_numpy(f["cidm_end"]).float().cuda()
    out = milnce_b200.SDTW_CIDM(Args())(v, t, start, end)
    out.backward()
    torch.cuda.synchronize()
    assert abs(out.item() - float(f["cidm_out"])) < 1e-5 * abs(float(f["cidm_out"]))
    assert rel_err(v.grad.cpu().numpy(), f["cidm_dvideo"]) < 1e-4
    assert rel_err(t.grad.cpu().numpy(), f["cidm_dtext"]) < 1e-4


def test_sdtw_negative_matches_executed_reference():
    """At the size the reference hard-codes (160 sequences x 8 frames x 512, loss.py:80-88)."""
    f = _fix()
    v = (torch.from_numpy(f["neg_video_i8"]).float() / 32).cuda().requires_grad_(True)
    t = (torch.from_numpy(f["neg_text_i8"]).float() / 32).cuda().requires_grad_(True)
    out = milnce_b200.SDTW_negative(Args())(v, t)
    out.backward()
    torch.cuda.synchronize()
    assert abs(out.item() - float(f["neg_out"])) < 1e-5 * abs(float(f["neg_out"]))
    assert rel_err(v.grad[:2].cpu().numpy(), f["neg_dvideo_first2"]) < 1e-4
    assert rel_err(t.grad[:2].cpu().numpy(), f["neg_dtext_first2"]) < 1e-4
    assert abs(v.grad.double().norm().item() - float(f["neg_dvideo_norm"])) < 1e-4 * float(f["neg_dvideo_norm"])
    assert abs(t.grad.double().norm().item() - float(f["neg_dtext_norm"])) < 1e-4 * float(f["neg_dtext_norm"])


def test_normalize_reads_operands_in_place():
    """normalize=True without the reference's cat([X, X, Y]) / cat([Y, X, Y]) copies: same value and
    gradients as the stacked formulation evaluated through this package."""
    g = torch.Generator().manual_seed(17)
    X = torch.randn(5, 20, 32, generator=g).cuda().requires_grad_(True)
    Y = torch.randn(5, 20, 32, generator=g).cuda().requires_grad_(True)
    out = SoftDTW(True, gamma=0.3, normalize=True, dist_func="cosine")(X, Y)
    out.sum().backward()
    dX, dY = X.grad.clone(), Y.grad.clone()
    X.grad = Y.grad = None
    plain = SoftDTW(True, gamma=0.3, dist_func="cosine")
    o3 = plain(torch.cat([X, X, Y]), torch.cat([Y, X, Y]))
    oxy, oxx, oyy = torch.split(o3, 5)
    ref = oxy - 0.5 * (oxx + oyy)
    ref.sum().backward()
    torch.cuda.synchronize()
    assert torch.allclose(out, ref, rtol=1e-6, atol=1e-5)
    assert rel_err(dX.cpu().numpy(), X.grad.cpu().numpy()) < 1e-5
    assert rel_err(dY.cpu().numpy(), Y.grad.cpu().numpy()) < 1e-5
