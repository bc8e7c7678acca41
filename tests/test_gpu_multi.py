"""Two-GPU NCCL check of GatheredMILNCELoss (skipped with fewer than two devices).

Each rank holds its shard of the global batch; the sharded module must return the same
loss on every rank and the local gradient slices of the single-GPU MILNCELoss evaluated
on the gathered tensors (what reference utils.py:20-24 hands back)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, scheme, out, gather="nccl", reduce="auto"):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import milnce_b200
        B, K, D = 1024, 4, 512
        b = B // world
        g = torch.Generator().manual_seed(77)
        V = (torch.randn(B, D, generator=g) * 0.1).bfloat16().cuda()
        T = (torch.randn(B * K, D, generator=g) * 0.1).bfloat16().cuda()
        v_loc = V[rank * b:(rank + 1) * b].clone().requires_grad_(True)
        t_loc = T[rank * b * K:(rank + 1) * b * K].clone().requires_grad_(True)
        mod = milnce_b200.GatheredMILNCELoss(scheme=scheme, gather=gather, reduce=reduce)
        loss = mod(v_loc, t_loc)
        loss.backward()
        if gather == "peer":
            assert mod.last_gather == "peer"
            # a second step through the same ring (slot reuse, barriers) must give the same result
            g1 = v_loc.grad.clone()
            v_loc.grad = None
            t_loc.grad = None
            loss = mod(v_loc, t_loc)
            loss.backward()
            assert torch.equal(g1, v_loc.grad)
        Vf = V.clone().requires_grad_(True)
        Tf = T.clone().requires_grad_(True)
        ref = milnce_b200.MILNCELoss()(Vf, Tf)
        ref.backward()
        torch.cuda.synchronize()
        ok = abs(loss.item() - ref.item()) < 1e-5 * abs(ref.item())
        dv_ref = Vf.grad[rank * b:(rank + 1) * b].float()
        dt_ref = Tf.grad[rank * b * K:(rank + 1) * b * K].float()
        e1 = ((v_loc.grad.float() - dv_ref).norm() / dv_ref.norm()).item()
        e2 = ((t_loc.grad.float() - dt_ref).norm() / dt_ref.norm()).item()
        ok = ok and e1 < 5e-3 and e2 < 5e-3   # both sides are rounded to bf16 by the module API
        out[rank] = 1 if ok else 0
        if not ok:
            print(f"rank {rank}: loss {loss.item()} vs {ref.item()} dV {e1} dT {e2}", flush=True)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("scheme", ["rows", "rowcol"])
def test_gathered_loss_two_gpus_nccl(scheme):
    world = 2
    out = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker, args=(world, _free_port(), scheme, out), nprocs=world, join=True)
    assert list(out) == [1] * world


def _worker_graphed(rank, world, port, out, gather="nccl"):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import milnce_b200
        B, K, D = 1024, 4, 512
        b = B // world
        step = milnce_b200.GraphedGatheredMILNCEStep(b, K, D, scheme="rows", partial_dtype=torch.bfloat16, gather=gather)
        ok = True
        for seed in (5, 6):     # two replays on different inputs
            g = torch.Generator().manual_seed(seed)
            V = (torch.randn(B, D, generator=g) * 0.1).bfloat16().cuda()
            T = (torch.randn(B * K, D, generator=g) * 0.1).bfloat16().cuda()
            loss, dv, dt = step(V[rank * b:(rank + 1) * b], T[rank * b * K:(rank + 1) * b * K])
            Vf = V.clone().requires_grad_(True)
            Tf = T.clone().requires_grad_(True)
            ref = milnce_b200.MILNCELoss()(Vf, Tf)
            ref.backward()
            torch.cuda.synchronize()
            dv_ref = Vf.grad[rank * b:(rank + 1) * b].float()
            dt_ref = Tf.grad[rank * b * K:(rank + 1) * b * K].float()
            e1 = ((dv.float() - dv_ref).norm() / dv_ref.norm()).item()
            e2 = ((dt.float() - dt_ref).norm() / dt_ref.norm()).item()
            good = abs(loss.item() - ref.item()) < 1e-5 * abs(ref.item()) and e1 < 5e-3 and e2 < 5e-3
            if not good:
                print(f"rank {rank} seed {seed}: loss {loss.item()} vs {ref.item()} dV {e1} dT {e2}", flush=True)
            ok = ok and good
        step.close()
        out[rank] = 1 if ok else 0
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.timeout(240)
def test_graphed_gathered_step_two_gpus_nccl():
    """The whole sharded step, NCCL collectives included, captured as one CUDA graph and replayed."""
    world = 2
    out = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker_graphed, args=(world, _free_port(), out), nprocs=world, join=True)
    assert list(out) == [1] * world


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.timeout(240)
@pytest.mark.parametrize("reduce", ["pull", "push", "nccl"])
def test_gathered_loss_peer_memory_two_gpus(reduce):
    """gather="peer": no embedding all-gather; the shards are published in symmetric memory and staged from the
    peers' buffers (milnce_peer_publish / milnce_forward_shard_peers).  reduce="push": the text-gradient GEMM
    delivers each owner's rows into that rank's memory (milnce_backward_shard_push) and every rank sums what it
    received (milnce_peer_reduce_dtext); "pull": partials stay local and are read from the peers; "nccl":
    reduce-scatter."""
    world = 2
    out = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker, args=(world, _free_port(), "rows", out, "peer", reduce), nprocs=world, join=True)
    assert list(out) == [1] * world


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.timeout(240)
def test_graphed_gathered_step_peer_memory_two_gpus():
    world = 2
    out = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker_graphed, args=(world, _free_port(), out, "peer"), nprocs=world, join=True)
    assert list(out) == [1] * world
