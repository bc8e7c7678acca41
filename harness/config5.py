#!/usr/bin/env python
"""BASELINE.json configs[4]: the S3D-G MIL-NCE train step with the loss head swapped.

    python harness/config5.py [--batch-per-gpu 32] [--steps 6] [--frames 32] [--size 224]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29544 harness/config5.py          # 8 x 32 = the global batch of 256

The towers are the REFERENCE's own model code (s3dg.py:207-328, imported unmodified from oracle/_ref - see
oracle/make_ref.py; random-init weights, synthetic clips and token ids, a synthetic dict.npy /
word2vec.pth of the real sizes, since the reference cannot construct S3D without those files).  What is
swapped is the three-line loss head of the train step (main_distributed.py:233-238):

    arm "reference":  allgather(video), allgather(text)  (utils.py:8-24)  ->  MILNCELoss (loss.py:6-18),
                      loss.item() every step (main_distributed.py:241)
    arm "b200":       GatheredMILNCELoss (all-gather + sharded loss + local gradient slices in one module),
                      the loss stays on the device and is read every `--display` steps

Both arms run the same model, the same optimiser and the same inputs under bf16 autocast.  Parity of the
head is checked on the embeddings the towers emit, against the reference's head evaluated in float32 on
those same values (under autocast the reference rounds its logits to bf16, which is reported beside it).
Prints one JSON line (rank 0): step time of both arms, the time of the loss head inside the step, its
share, and the parity errors.  TEST / BENCH INFRASTRUCTURE: this file uses oracle/_ref.
"""
from __future__ import annotations

import argparse
import copy
import json
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def synthetic_assets(tmp):
    """dict.npy (66 250 tokens) and word2vec.pth (66 250 x 300): the files S3D() loads at construction."""
    d = os.path.join(tmp, "dict.npy")
    w = os.path.join(tmp, "word2vec.pth")
    np.save(d, np.array([f"w{i}" for i in range(66250)]))
    g = torch.Generator().manual_seed(7)
    torch.save(torch.randn(66250, 300, generator=g) * 0.1, w)
    return d, w


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch-per-gpu", type=int, default=32)
    ap.add_argument("--candidates", type=int, default=4)
    ap.add_argument("--frames", type=int, default=32)
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--display", type=int, default=10, help="the b200 arm reads the loss every this many steps")
    args = ap.parse_args()

    from oracle import ref_loader
    why = ref_loader.why_unavailable()
    if why or not os.path.exists(os.path.join(ref_loader.REF_DIR, "s3dg.py")):
        print(json.dumps({"config": 5, "unavailable": why or "oracle/_ref/s3dg.py missing (python oracle/make_ref.py)"}))
        return
    import milnce_b200

    # the JSON line must be the only thing on stdout, and NCCL prints its version banner there
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    class A:          # what the reference's args carries into AllGather / the train step
        pass
    a = A()
    a.world_size, a.rank, a.distributed, a.gpu = world, rank, world > 1, local

    s3dg = ref_loader.load("s3dg")
    ref_loss = ref_loader.load("loss")
    ref_utils = ref_loader.load("utils")
    with tempfile.TemporaryDirectory() as tmp:
        dict_path, w2v_path = synthetic_assets(tmp)
        torch.manual_seed(0)
        model0 = s3dg.S3D(num_classes=512, word2vec_path=w2v_path, token_to_word_path=dict_path)
    model0 = model0.to(dev)

    b, K = args.batch_per_gpu, args.candidates
    g = torch.Generator().manual_seed(100 + rank)
    video = torch.randint(0, 256, (b, 3, args.frames, args.size, args.size), dtype=torch.uint8, generator=g).to(dev)
    text = torch.randint(0, 66250, (b, K, 16), generator=g).to(dev)

    def towers(model):
        v = video.float() / 255.0                                  # main_distributed.py:229-231
        t = text.view(-1, text.shape[-1])
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return model(v, t)

    allgather = ref_utils.AllGather.apply
    ref_head = ref_loss.MILNCELoss()
    our_head = milnce_b200.GatheredMILNCELoss(args=a if world > 1 else None)

    def head_reference(ve, te):
        if world > 1:
            ve = allgather(ve, a)
            te = allgather(te, a)
        return ref_head(ve, te)

    def head_b200(ve, te):
        return our_head(ve, te)

    def run_arm(head, read_every):
        model = copy.deepcopy(model0)
        opt = torch.optim.Adam(model.parameters(), lr=1e-4)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        times, head_times = [], []
        first = None
        losses = []
        for it in range(args.warmup + args.steps):
            ev[0].record()
            opt.zero_grad()
            ve, te = towers(model)
            if it == 0:
                ve.retain_grad()
                te.retain_grad()
            ev[1].record()
            loss = head(ve, te)
            ev[2].record()
            loss.backward()
            opt.step()
            losses.append(loss.detach())
            if (it + 1) % read_every == 0:
                _ = float(torch.stack(losses).sum().item())       # the host reads the loss here
                losses = []
            ev[3].record()
            torch.cuda.synchronize()
            if it == 0:
                first = (loss.item(), ve.grad.float().clone(), te.grad.float().clone())
            if it >= args.warmup:
                times.append(ev[0].elapsed_time(ev[3]))
                head_times.append(ev[1].elapsed_time(ev[2]))
        del model, opt
        torch.cuda.empty_cache()
        return sum(times) / len(times), sum(head_times) / len(head_times), first

    def rel(x, y):
        return ((x.double() - y.double()).norm() / y.double().norm().clamp_min(1e-30)).item()

    # Parity of the loss head on the embeddings the towers emit (bf16 under autocast).  The yardstick is the
    # reference's own head evaluated in float32 on those same bf16 values; the reference arm AS RUN (autocast:
    # its matmul rounds the logits to bf16, loss.py:11) is reported next to it - its distance to that yardstick
    # is the noise floor of the comparison between the two arms' first steps below.
    with torch.no_grad():
        ve0, te0 = towers(model0)

    def head_grads(head, ve, te, autocast):
        ve = ve.detach().clone().requires_grad_(True)
        te = te.detach().clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            loss = head(ve, te)
        loss.backward()
        return loss.item(), ve.grad.float(), te.grad.float()

    y_loss, y_dv, y_dt = head_grads(head_reference, ve0.float(), te0.float(), False)
    o_loss, o_dv, o_dt = head_grads(head_b200, ve0, te0, True)
    r_loss, r_dv, r_dt = head_grads(head_reference, ve0, te0, True)
    head_parity = {
        "b200_vs_reference_head_in_float32": {"loss_rel": abs(o_loss - y_loss) / abs(y_loss),
                                              "dvideo_embd_rel": rel(o_dv, y_dv), "dtext_embd_rel": rel(o_dt, y_dt)},
        "reference_under_autocast_vs_same": {"loss_rel": abs(r_loss - y_loss) / abs(y_loss),
                                             "dvideo_embd_rel": rel(r_dv, y_dv), "dtext_embd_rel": rel(r_dt, y_dt)},
        "tolerance": {"loss_rel": 1e-5, "grad_rel": 5e-3,
                      "note": "gradients are returned in the embeddings' dtype (bf16): one rounding, 2^-9 per element"},
    }
    hp = head_parity["b200_vs_reference_head_in_float32"]
    ok = hp["loss_rel"] < 1e-5 and hp["dvideo_embd_rel"] < 5e-3 and hp["dtext_embd_rel"] < 5e-3

    t_ref, h_ref, f_ref = run_arm(head_reference, 1)
    t_our, h_our, f_our = run_arm(head_b200, args.display)

    out = {
        "config": 5, "workload": f"S3D-G MIL-NCE train step, {args.frames}x{args.size}x{args.size} clips, "
                                 f"{b} per GPU x {world} GPUs = global batch {b * world}, K={K}, bf16 autocast, Adam",
        "n_gpus": world, "data": "synthetic clips / token ids, random-init reference S3D (oracle/_ref/s3dg.py)",
        "reference_arm": {"ms_per_step": t_ref, "loss_head_forward_ms": h_ref,
                          "head": "AllGather x 2 + MILNCELoss (reference), loss.item() every step"},
        "b200_arm": {"ms_per_step": t_our, "loss_head_forward_ms": h_our,
                     "head": f"GatheredMILNCELoss, loss read every {args.display} steps"},
        "step_speedup": t_ref / t_our, "loss_head_forward_share_reference": h_ref / t_ref,
        "loss_head_forward_share_b200": h_our / t_our,
        "head_parity": head_parity,
        "first_step_arm_vs_arm": {"loss_rel": abs(f_our[0] - f_ref[0]) / abs(f_ref[0]),
                                  "dvideo_embd_rel": rel(f_our[1], f_ref[1]), "dtext_embd_rel": rel(f_our[2], f_ref[2]),
                                  "note": "informational: the reference arm rounds its logits to bf16 under autocast"},
    }
    out["parity_ok"] = bool(ok)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank == 0:
        print(json.dumps(out), flush=True)
    os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        sys.exit(3)


if __name__ == "__main__":
    main()
