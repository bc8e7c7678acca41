"""The CPU oracle against fixtures minted from the executed reference
(oracle/make_golden.py).  Runs without a GPU."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err
from oracle import milnce_oracle, sdtw_oracle

MILNCE_FILES = sorted(glob.glob(os.path.join(GOLDEN, "milnce_*.npz")))
SDTW_FILES = sorted(glob.glob(os.path.join(GOLDEN, "sdtw_*.npz")))


def test_fixtures_present():
    assert len(MILNCE_FILES) >= 5 and len(SDTW_FILES) >= 6


@pytest.mark.parametrize("path", MILNCE_FILES, ids=os.path.basename)
def test_milnce_closed_form_matches_reference(path):
    g = np.load(path)
    loss, dv, dt, den, nom = milnce_oracle.milnce_closed_form(g["video"], g["text"])
    tol = 1e-6 if g["dvideo"].dtype == np.float32 else 1e-12
    assert abs(loss - float(g["loss"])) <= 1e-12 * max(1.0, abs(float(g["loss"])))
    assert rel_err(dv, g["dvideo"]) < tol
    assert rel_err(dt, g["dtext"]) < tol
    assert np.all(den >= nom - 1e-12)


@pytest.mark.parametrize("path", MILNCE_FILES, ids=os.path.basename)
def test_milnce_torch_ports_match_reference(path):
    import torch
    g = np.load(path)
    for fn in (milnce_oracle.milnce_torch_port, milnce_oracle.milnce_torch_lean):
        v = torch.from_numpy(g["video"]).double().requires_grad_(True)
        t = torch.from_numpy(g["text"]).double().requires_grad_(True)
        out = fn(v, t)
        out.backward()
        assert abs(out.item() - float(g["loss"])) < 1e-11 * max(1.0, abs(float(g["loss"])))
        tol = 1e-6 if g["dvideo"].dtype == np.float32 else 1e-11
        assert rel_err(v.grad.numpy(), g["dvideo"]) < tol
        assert rel_err(t.grad.numpy(), g["dtext"]) < tol


def _keys(g):
    return sorted(k[:-4] for k in g.files if k.endswith("_out") and k != "grad_out")


@pytest.mark.parametrize("path", SDTW_FILES, ids=os.path.basename)
def test_sdtw_recurrences_match_reference_f64(path):
    g = np.load(path)
    bw = float(g["bandwidth"])
    for key in _keys(g):
        if key + "_R" not in g.files:
            continue
        dist, gam, _ = key.split("_g")[0], float(key.split("_g")[1].split("_n")[0]), None
        gamma = float(np.float32(gam))
        D = sdtw_oracle.dist_forward(g["x"], g["y"], dist)
        assert rel_err(D, g[key + "_D"]) < 1e-13
        R = sdtw_oracle.dp_forward(g[key + "_D"], gamma, bw)
        Rg = g[key + "_R"]
        assert np.array_equal(np.isinf(R), np.isinf(Rg))
        fin = np.isfinite(Rg)
        assert np.max(np.abs(R[fin] - Rg[fin])) <= 1e-12 * max(1.0, np.max(np.abs(Rg[fin])))
        E = sdtw_oracle.dp_backward(g[key + "_D"], R, gamma, bw)
        assert rel_err(E, g[key + "_E"]) < 1e-12
        # C code against the pure-Python loops
        Rp = sdtw_oracle.dp_forward_py(g[key + "_D"][:1], gamma, bw)
        assert np.allclose(Rp[np.isfinite(Rp)], R[:1][np.isfinite(Rp)], rtol=1e-13, atol=1e-13)


@pytest.mark.parametrize("path", SDTW_FILES, ids=os.path.basename)
def test_sdtw_module_matches_reference(path):
    """End-to-end SoftDTW(use_cuda=False).  The reference truncates R and E to
    float32 between its numba kernels and autograd (soft_dtw_cuda.py:257,270), so
    agreement is at float32 resolution of R divided by gamma."""
    g = np.load(path)
    bw = float(g["bandwidth"])
    for key in _keys(g):
        dist = key.split("_g")[0]
        gam = float(key.split("_g")[1].split("_n")[0])
        normalize = key.endswith("_n1")
        out, dx, dy, _ = sdtw_oracle.softdtw(g["x"], g["y"], gamma=gam, normalize=normalize,
                                             bandwidth=(bw if bw > 0 else None), dist_func=dist,
                                             grad_out=g["grad_out"])
        scale = np.max(np.abs(g[key + "_out"])) + 1.0
        assert np.max(np.abs(out - g[key + "_out"])) < 5e-6 * scale, key
        # float32 truncation of R (|R| ~ scale) enters the E recurrence divided by gamma
        # (the normalised variant subtracts three such terms, so its noise is larger)
        gtol = max(2e-5, (64 if normalize else 8) * 2.0 ** -24 * scale / gam)
        if dist == "euclidean" and normalize:
            # x-vs-x distances are zero on the diagonal: sqrt backward gives NaN in
            # the reference as well; compare NaN patterns only.
            assert np.array_equal(np.isnan(dx), np.isnan(g[key + "_dx"])), key
            continue
        assert rel_err(dx, g[key + "_dx"]) < gtol, key
        assert rel_err(dy, g[key + "_dy"]) < gtol, key


def test_sdtw_negative_restatement_matches_reference():
    """The (b, L, b, L)-view form of reference loss.py:77-91 that milnce_b200.SDTW_negative uses (diagonal blocks
    set to 0 before the exp, divisor b - 1), evaluated in float64 with the oracle's soft-DTW, against the executed
    reference at the size it hard-codes (160 x 8 x 512)."""
    import os

    import numpy as np

    from oracle import sdtw_oracle
    f = np.load(os.path.join(os.path.dirname(__file__), "golden", "composite_losses.npz"))
    v = f["neg_video_i8"].astype(np.float64) / 32
    t = f["neg_text_i8"].astype(np.float64) / 32
    b, L, d = v.shape
    sdtw, dv_s, dt_s, _ = sdtw_oracle.softdtw(v, t, gamma=0.1, dist_func="cosine", grad_out=np.full(b, 1.0 / b))
    pw = (v.reshape(-1, d) @ t.reshape(-1, d).T).reshape(b, L, b, L)
    for i in range(b):
        pw[i, :, i, :] = 0.0
    e = np.exp(pw)
    neg = e.sum(axis=(1, 2, 3))
    loss = np.mean(sdtw + neg / (b - 1))
    assert abs(loss - float(f["neg_out"])) < 1e-9 * abs(float(f["neg_out"]))
    # gradient of the negative term: masked entries are constants
    for i in range(b):
        e[i, :, i, :] = 0.0
    w = e / (b * (b - 1))
    dv = dv_s + np.einsum("ipjq,jqd->ipd", w, t)
    dt = dt_s + np.einsum("ipjq,ipd->jqd", w, v)
    # (the reference CPU soft-DTW hands R back through float32, soft_dtw_cuda.py:257, and its backward divides
    # differences of R by gamma = 0.1: a few 1e-6 on the gradients)
    def rel(a, r):
        return np.linalg.norm(a - r) / np.linalg.norm(r)
    assert rel(dv[:2], f["neg_dvideo_first2"]) < 1e-5
    assert rel(dt[:2], f["neg_dtext_first2"]) < 1e-5
