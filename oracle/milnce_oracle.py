"""CPU oracle for the MIL-NCE loss head.  TEST INFRASTRUCTURE ONLY.

This file is a checker: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  The
product path (``mil-nce_howto100m_b200``) never does.

Two restatements of the reference's ``MILNCELoss`` (reference ``loss.py:10-18``):

* :func:`milnce_closed_form` - float64 numpy closed form of loss and gradients
  (SURVEY.md appendix A).  Pinned against the executed reference by
  ``oracle/make_golden.py`` -> ``tests/golden/milnce_*.npz``.
* :func:`milnce_torch_port` - a torch port that keeps the reference's operator
  sequence (matmul, identity mask, sum, logsumexp, concatenation, logsumexp,
  mean) so that timing it on host cores is a fair stand-in for the reference's
  CPU path (the reference itself cannot run on CPU: ``loss.py:13`` calls
  ``.cuda()`` on the mask).

Layout (reference ``train.py:212-213``): ``video`` is (B, D); ``text`` is
(B*K, D), sample-major, row ``j*K+k`` = candidate ``k`` of sample ``j``.
"""
from __future__ import annotations

import numpy as np


def _lse(a: np.ndarray, axis) -> np.ndarray:
    m = np.max(a, axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    out = np.log(np.sum(np.exp(a - m), axis=axis, keepdims=True)) + m
    return np.squeeze(out, axis=axis)


def milnce_stats(video: np.ndarray, text: np.ndarray):
    """Return (x, nom, den) in float64.

    x[i, j, k] = <video_i, text_{jK+k}>                       (loss.py:11-12)
    nom[i]     = logsumexp_k x[i, i, k]                       (loss.py:13-15)
    den[i]     = logsumexp over x[i, :, :] and x[:, i, :]     (loss.py:16-17)
    """
    v = np.asarray(video, dtype=np.float64)
    t = np.asarray(text, dtype=np.float64)
    b = v.shape[0]
    if b == 0 or t.shape[0] % b != 0:
        raise ValueError("text rows must be a multiple of video rows")
    k = t.shape[0] // b
    x = (v @ t.T).reshape(b, b, k)
    idx = np.arange(b)
    nom = _lse(x[idx, idx, :], axis=1)
    row = _lse(x.reshape(b, b * k), axis=1)
    col = _lse(np.transpose(x, (1, 0, 2)).reshape(b, b * k), axis=1)
    den = np.logaddexp(row, col)
    return x, nom, den


def milnce_closed_form(video: np.ndarray, text: np.ndarray, grad_loss: float = 1.0):
    """float64 loss, d loss/d video, d loss/d text, den, nom.

    loss = mean(den - nom)                                    (loss.py:18)
    G[i,j,k] = (e^{x-den_i} + e^{x-den_j} - [i==j] e^{x-nom_i}) / B
    dV = G . T ; dT = G^T . V          (autograd of loss.py:11-18, closed form)
    """
    v = np.asarray(video, dtype=np.float64)
    t = np.asarray(text, dtype=np.float64)
    x, nom, den = milnce_stats(v, t)
    b, _, k = x.shape
    loss = float(np.mean(den - nom))
    g = np.exp(x - den[:, None, None]) + np.exp(x - den[None, :, None])
    idx = np.arange(b)
    g[idx, idx, :] -= np.exp(x[idx, idx, :] - nom[:, None])
    g *= grad_loss / b
    g2 = g.reshape(b, b * k)
    dv = g2 @ t
    dt = g2.T @ v
    return loss, dv, dt, den, nom


def milnce_torch_port(video, text):
    """Torch port with the reference's operator sequence (loss.py:10-18).

    Differentiable through autograd, runs on whatever device the inputs are on.
    Used (a) as the timed CPU baseline and (b) as a float64 checker on large
    shapes where numpy would be slow.
    """
    import torch

    b = video.shape[0]
    sim = (video @ text.transpose(0, 1)).reshape(b, b, -1)
    mask = torch.eye(b, dtype=sim.dtype, device=sim.device).unsqueeze(-1)
    positives = (sim * mask).sum(dim=1)
    nom = torch.logsumexp(positives, dim=1)
    both = torch.cat((sim, sim.transpose(0, 1)), dim=1).reshape(b, -1)
    den = torch.logsumexp(both, dim=1)
    return (den - nom).mean()


def milnce_torch_lean(video, text):
    """Memory-lean float64-capable checker for large B (no B x 2BK concat).

    Same value as :func:`milnce_torch_port`; used by GPU tests at B=8192 where
    the concat of the port would need several GB in float64.
    """
    import torch

    b = video.shape[0]
    sim = (video @ text.transpose(0, 1)).reshape(b, b, -1)
    idx = torch.arange(b, device=sim.device)
    nom = torch.logsumexp(sim[idx, idx, :], dim=1)
    row = torch.logsumexp(sim.reshape(b, -1), dim=1)
    col = torch.logsumexp(sim.transpose(0, 1).reshape(b, -1), dim=1)
    den = torch.logaddexp(row, col)
    return (den - nom).mean()


def sharded_slices(b: int, k: int, world: int, rank: int):
    """Row ranges a rank owns under the reference's AllGather (utils.py:11-24).

    Returns ((v0, v1), (t0, t1)): rows of the gathered video / text tensors that
    ``AllGather.backward`` hands back to ``rank``.
    """
    if b % world:
        raise ValueError("global batch must divide by world size")
    per = b // world
    return (per * rank, per * (rank + 1)), (per * k * rank, per * k * (rank + 1))
