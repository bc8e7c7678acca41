"""World-size-2 (and 4) checks of the multi-GPU host logic on CPU with gloo.

The CUDA kernels cannot run here, so the range-restricted compute backend is
replaced by the CPU oracle (tests may use it as a stand-in).  What is verified is
the orchestration that bench.py --gpus N and the train-step call site rely on:
* ``AllGather`` reproduces reference utils.py:11-24 (rank-ordered concat forward,
  local slice backward with no reduction);
* ``GatheredMILNCELoss`` shards the loss over ranks (row block + column block per
  rank, one all-gather of den and the loss shares) and still returns the global
  loss and exactly the local gradient slices of the gathered reference loss.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


class OracleBackend:
    """CPU stand-in with the C-ABI's range semantics (include/b200head.h)."""

    @staticmethod
    def forward(V, T, i0, n):
        from oracle import milnce_oracle
        B = V.shape[0]
        _, nom, den = milnce_oracle.milnce_stats(V.double().numpy(), T.double().numpy())
        den_out = np.full(B, np.nan)
        nom_out = np.full(B, np.nan)
        den_out[i0:i0 + n] = den[i0:i0 + n]   # a rank only knows its own entries
        nom_out[i0:i0 + n] = nom[i0:i0 + n]
        part = np.sum(den[i0:i0 + n] - nom[i0:i0 + n]) / B
        return (torch.tensor([part], dtype=torch.float64), torch.from_numpy(den_out),
                torch.from_numpy(nom_out), None)

    @staticmethod
    def backward(V, T, den, nom, g, i0, n, ws):
        v = V.double().numpy()
        t = T.double().numpy()
        B = v.shape[0]
        K = t.shape[0] // B
        den = den.numpy()
        nom = nom.numpy()
        assert not np.isnan(den).any(), "den must be all-gathered before backward"
        x = (v @ t.T).reshape(B, B, K)
        scale = float(g[0]) / B
        rows = slice(i0, i0 + n)
        # row block: G[rows, :, :]
        g_r = np.exp(x[rows] - den[rows, None, None]) + np.exp(x[rows] - den[None, :, None])
        for r in range(n):
            g_r[r, i0 + r, :] -= np.exp(x[i0 + r, i0 + r, :] - nom[i0 + r])
        dv = scale * g_r.reshape(n, B * K) @ t
        # column block: G[:, rows, :]
        g_c = np.exp(x[:, rows] - den[:, None, None]) + np.exp(x[:, rows] - den[None, rows, None])
        for r in range(n):
            g_c[i0 + r, r, :] -= np.exp(x[i0 + r, i0 + r, :] - nom[i0 + r])
        dt = scale * g_c.reshape(B, n * K).T @ v
        return torch.from_numpy(dv), torch.from_numpy(dt)


    # ---- rows-only sharding (include/b200head.h: milnce_forward_shard / milnce_backward_shard) ----
    @staticmethod
    def supports_rows_scheme(video, i0):
        return True

    @staticmethod
    def forward_shard(V, T, i0, n):
        from oracle import milnce_oracle
        B = V.shape[0]
        x, nom, _ = milnce_oracle.milnce_stats(V.double().numpy(), T.double().numpy())
        K = x.shape[2]
        row = np.full(B, np.nan)
        nm = np.full(B, np.nan)
        row[i0:i0 + n] = milnce_oracle._lse(x[i0:i0 + n].reshape(n, B * K), axis=1)
        nm[i0:i0 + n] = nom[i0:i0 + n]
        # partial column statistic over the OWNED rows only, for every sample j
        colp = milnce_oracle._lse(np.transpose(x[i0:i0 + n], (1, 0, 2)).reshape(B, n * K), axis=1)
        return torch.from_numpy(np.concatenate([row[i0:i0 + n], nm[i0:i0 + n], colp])), None

    @staticmethod
    def shard_combine(allp, world, b):
        B = world * b
        row_all = allp[:, :b].reshape(B)
        nom_all = allp[:, b:2 * b].reshape(B).contiguous()
        col_all = torch.logsumexp(allp[:, 2 * b:], dim=0)
        den = torch.logaddexp(row_all, col_all).contiguous()
        return den, nom_all, (den - nom_all).mean()

    @staticmethod
    def backward_shard(V, T, den, nom, g, i0, n, ws, parts=3, grad_dtype=None, partial_dtype=None, dims=None):
        v = V.double().numpy()
        t = T.double().numpy()
        B = v.shape[0]
        K = t.shape[0] // B
        den = den.numpy()
        nom = nom.numpy()
        assert not np.isnan(den).any()
        x = (v @ t.T).reshape(B, B, K)
        rows = slice(i0, i0 + n)
        g_r = np.exp(x[rows] - den[rows, None, None]) + np.exp(x[rows] - den[None, :, None])
        for r in range(n):
            g_r[r, i0 + r, :] -= np.exp(x[i0 + r, i0 + r, :] - nom[i0 + r])
        g_r = (float(g[0]) / B) * g_r.reshape(n, B * K)
        dv = torch.from_numpy(g_r @ t) if parts & 2 else None
        dtp = torch.from_numpy(g_r.T @ v[rows]) if parts & 1 else None
        return dv, dtp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, b, K, D, scheme, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import milnce_b200
        from oracle import milnce_oracle

        class Args:
            pass
        args = Args()
        args.world_size, args.rank = world, rank
        g = torch.Generator().manual_seed(1234)
        V = torch.randn(world * b, D, generator=g, dtype=torch.float64) * 0.3
        T = torch.randn(world * b * K, D, generator=g, dtype=torch.float64) * 0.3
        v_loc = V[rank * b:(rank + 1) * b].clone().requires_grad_(True)
        t_loc = T[rank * b * K:(rank + 1) * b * K].clone().requires_grad_(True)

        # 1. AllGather semantics (reference utils.py:11-24)
        gv = milnce_b200.AllGather.apply(v_loc, args)
        assert torch.equal(gv.detach(), V)
        w = torch.arange(gv.numel(), dtype=torch.float64).reshape(gv.shape)
        (gv * w).sum().backward()
        assert torch.equal(v_loc.grad, w[rank * b:(rank + 1) * b])  # local slice, no reduction
        v_loc.grad = None

        # 2. sharded loss == reference loss on gathered tensors, grads == local slices
        mod = milnce_b200.GatheredMILNCELoss(args=args, _backend=OracleBackend, scheme=scheme)
        loss = mod(v_loc, t_loc)
        (loss * 1.75).backward()  # 1.75 is exact in float32 (grad_loss is passed as fp32)
        ref_loss, ref_dv, ref_dt, _, _ = milnce_oracle.milnce_closed_form(V.numpy(), T.numpy(), 1.75)
        ok = abs(loss.item() - ref_loss) < 1e-12
        ok &= np.allclose(v_loc.grad.numpy(), ref_dv[rank * b:(rank + 1) * b], rtol=1e-11, atol=1e-14)
        ok &= np.allclose(t_loc.grad.numpy(), ref_dt[rank * b * K:(rank + 1) * b * K], rtol=1e-11,
                          atol=1e-14)
        out[rank] = 1 if ok else 0
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("scheme", ["rowcol", "rows"])
@pytest.mark.parametrize("world", [2, 4])
def test_gathered_loss_world(world, scheme):
    port = _free_port()
    out = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker, args=(world, port, 3, 2, 8, scheme, out), nprocs=world, join=True)
    assert list(out) == [1] * world


def test_world_size_one_needs_no_process_group():
    sys.path.insert(0, ROOT)
    import milnce_b200
    from oracle import milnce_oracle
    g = torch.Generator().manual_seed(5)
    V = (torch.randn(6, 8, generator=g, dtype=torch.float64)).requires_grad_(True)
    T = (torch.randn(12, 8, generator=g, dtype=torch.float64)).requires_grad_(True)
    loss = milnce_b200.GatheredMILNCELoss(_backend=OracleBackend)(V, T)
    loss.backward()
    ref_loss, ref_dv, ref_dt, _, _ = milnce_oracle.milnce_closed_form(V.detach().numpy(),
                                                                     T.detach().numpy())
    assert abs(loss.item() - ref_loss) < 1e-12
    assert np.allclose(V.grad.numpy(), ref_dv, rtol=1e-11, atol=1e-14)
    assert np.allclose(T.grad.numpy(), ref_dt, rtol=1e-11, atol=1e-14)


def test_option_names_are_validated():
    sys.path.insert(0, ROOT)
    import milnce_b200
    for kw in ({"scheme": "columns"}, {"gather": "mpi"}, {"reduce": "ring"}):
        with pytest.raises(ValueError):
            milnce_b200.GatheredMILNCELoss(**kw)
    # the documented values construct without a process group or a GPU
    for red in ("auto", "nccl", "pull", "push"):
        milnce_b200.GatheredMILNCELoss(gather="auto", reduce=red)
