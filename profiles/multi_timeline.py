"""Per-phase timeline of one GatheredMILNCELoss step on W GPUs (CUDA events, max over ranks).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
        --master-port 29511 profiles/multi_timeline.py [--out profiles/r2_multi_timeline_W.txt]

For every scheme ("rows" with float32 and with bfloat16 text-gradient partials, "rowcol") the step is
run with phase stamps (the module's _marks hook: one CUDA event on the compute stream after each
phase; a collective's stamp is recorded after its wait(), i.e. when the compute stream may go on).
Also timed alone: the collectives of the step at their real sizes.  B=8192, K=4, D=512, bf16.
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--only", default=None, help="run only the variants whose name contains this")
    args = ap.parse_args()
    import milnce_b200
    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    B, K, D = args.batch, 4, 512
    b = B // world
    g = torch.Generator().manual_seed(1000 + rank)
    v = (torch.randn(b, D, generator=g) * 0.05).bfloat16().to(dev).requires_grad_(True)
    t = (torch.randn(b * K, D, generator=g) * 0.05).bfloat16().to(dev).requires_grad_(True)
    lines = [f"# GatheredMILNCELoss step timeline, W={world}, B={B} K={K} D={D} bf16; microseconds, median over "
             f"{args.steps} steps, max over ranks"]

    def allmax(x):
        tt = torch.tensor(x, dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return tt.tolist()

    variants = [("rows/f32-partial", dict(scheme="rows", partial_dtype=torch.float32, gather="nccl")),
                ("rows/bf16-partial", dict(scheme="rows", partial_dtype=torch.bfloat16, gather="nccl")),
                ("rows/bf16-partial/peer-memory gather, nccl reduce-scatter",
                 dict(scheme="rows", partial_dtype=torch.bfloat16, gather="peer", reduce="nccl")),
                ("rows/bf16-partial/peer-memory gather, pull reduce from peer memory",
                 dict(scheme="rows", partial_dtype=torch.bfloat16, gather="peer", reduce="pull")),
                ("rows/bf16-partial/peer-memory gather, push from the dtext GEMM epilogue",
                 dict(scheme="rows", partial_dtype=torch.bfloat16, gather="peer", reduce="push")),
                ("rowcol", dict(scheme="rowcol"))]
    if args.only:
        variants = [v for v in variants if args.only in v[0]]
    for name, kw in variants:
        mod = milnce_b200.GatheredMILNCELoss(**kw)
        for _ in range(5):
            v.grad = t.grad = None
            mod(v, t).backward()
        # whole step, no stamps
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            v.grad = t.grad = None
            mod(v, t).backward()
        e1.record()
        torch.cuda.synchronize()
        whole = allmax([e0.elapsed_time(e1) / args.steps * 1e3])[0]
        # stamped steps
        per_phase = {}
        order = []
        for _ in range(args.steps):
            mod._marks = []
            v.grad = t.grad = None
            dist.barrier()
            mod(v, t).backward()
            torch.cuda.synchronize()
            mk = mod._marks
            for (n0, a), (n1, c) in zip(mk[:-1], mk[1:]):
                key = n1
                if key not in per_phase:
                    per_phase[key] = []
                    order.append(key)
                per_phase[key].append(a.elapsed_time(c) * 1e3)
        mod._marks = None
        med = [sorted(per_phase[k])[len(per_phase[k]) // 2] for k in order]
        med = allmax(med)
        # the same step captured as one CUDA graph (GraphedGatheredMILNCEStep) and replayed
        gs = milnce_b200.GraphedGatheredMILNCEStep(b, K, D, **kw)
        with torch.no_grad():
            gs.video.copy_(v)
            gs.text.copy_(t)
        for _ in range(3):
            gs.replay()
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            gs.replay()
        e1.record()
        torch.cuda.synchronize()
        graphed = allmax([e0.elapsed_time(e1) / args.steps * 1e3])[0]
        gs.close()
        lines.append(f"\n## {name}: whole step {whole:.0f} us eager, {graphed:.0f} us replayed as one CUDA graph "
                     f"(back to back, no stamps)")
        for k, m in zip(order, med):
            note = "   (autograd glue between forward and backward; host-side)" if k == "backward_start" else ""
            lines.append(f"  {k:34s} {m:8.0f}{note}")
        lines.append(f"  {'sum of phases':34s} {sum(med):8.0f}")

    # the collectives alone
    lines.append("\n## collectives alone (CUDA events around 20 back-to-back calls)")

    def time_coll(fn, n=20):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return allmax([e0.elapsed_time(e1) / n * 1e3])[0]

    Vb = torch.empty(B, D, dtype=torch.bfloat16, device=dev)
    Tb = torch.empty(B * K, D, dtype=torch.bfloat16, device=dev)
    vd, td = v.detach(), t.detach()

    def gather2():
        with dist._coalescing_manager(device=dev, async_ops=True) as cm:
            dist.all_gather_into_tensor(Vb, vd)
            dist.all_gather_into_tensor(Tb, td)
        cm.wait()

    def gather_sep():
        dist.all_gather_into_tensor(Vb, vd)
        dist.all_gather_into_tensor(Tb, td)

    small = torch.empty(2 * b + B, device=dev)
    small_all = torch.empty(world * (2 * b + B), device=dev)
    p32 = torch.empty(B * K, D, device=dev)
    o32 = torch.empty(b * K, D, device=dev)
    p16 = torch.empty(B * K, D, dtype=torch.bfloat16, device=dev)
    o16 = torch.empty(b * K, D, dtype=torch.bfloat16, device=dev)
    lines.append(f"  all-gather video+text, one grouped call ({(Vb.numel() + Tb.numel()) * 2 / 1e6:.0f} MB out)  {time_coll(gather2):8.0f}")
    lines.append(f"  all-gather video, text as two calls                          {time_coll(gather_sep):8.0f}")
    lines.append(f"  all-gather of the statistics ({small.numel() * 4 / 1e3:.0f} KB per rank)                {time_coll(lambda: dist.all_gather_into_tensor(small_all, small)):8.0f}")
    lines.append(f"  reduce-scatter dtext partial float32 ({p32.numel() * 4 / 1e6:.0f} MB in)              {time_coll(lambda: dist.reduce_scatter_tensor(o32, p32)):8.0f}")
    lines.append(f"  reduce-scatter dtext partial bfloat16 ({p16.numel() * 2 / 1e6:.0f} MB in)             {time_coll(lambda: dist.reduce_scatter_tensor(o16, p16)):8.0f}")
    # the peer-memory kernels alone (GPU time without host gaps: 20 back-to-back launches between two events)
    try:
        import ctypes
        from milnce_b200 import _lib
        from milnce_b200.utils import PeerRing
        lib = _lib.load()
        ring = PeerRing(b, K, D, dev)
        st = _lib.stream_ptr(dev)
        P = lambda x: ctypes.c_void_p(x.data_ptr())
        nom = torch.empty(b, device=dev)
        nbytes = lib.milnce_workspace_bytes(B, K, D, _lib.BF16, b)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        row = torch.empty(B, device=dev)
        slots = (ctypes.c_void_p * world)(*ring.ptrs)
        packs = (ctypes.c_void_p * world)(*ring.pack_ptrs)
        grads = (ctypes.c_void_p * world)(*ring.grad_ptrs)
        den, nm, ls = torch.empty(B, device=dev), torch.empty(B, device=dev), torch.empty(1, device=dev)
        dt16 = torch.empty(b * K, D, dtype=torch.bfloat16, device=dev)
        pub = lambda: lib.milnce_peer_publish(P(vd), P(td), b, K, D, ctypes.c_void_p(ring.ptrs[rank]), P(nom), st)
        pub()
        ring.barrier()
        fwd = lambda: lib.milnce_forward_shard_peers(slots, world, B, K, D, rank * b, b,
                                                     ctypes.c_void_p(row.data_ptr()), ctypes.c_void_p(ring.pack.data_ptr() + 8 * b),
                                                     P(ws), nbytes, st)
        assert fwd() == 0
        ring.barrier()
        lines.append("\n## peer-memory kernels alone (GPU time, 20 back-to-back launches; us)")
        lines.append(f"  symmetric-memory barrier                                      {time_coll(ring.barrier):8.0f}")
        lines.append(f"  milnce_peer_publish (header memset + copy of the shard + nom)   {time_coll(pub):8.0f}")
        lines.append(f"  milnce_forward_shard_peers (stage from W slots + similarity pass + merges) {time_coll(fwd):8.0f}")
        loc = torch.empty(B * K * D + B * D, dtype=torch.bfloat16, device=dev)
        lines.append(f"  milnce_shard_combine_peers                                    "
                     f"{time_coll(lambda: lib.milnce_shard_combine_peers(packs, world, b, P(den), P(nm), P(ls), P(ring.combine_scratch), st)):8.0f}")
        lines.append(f"  milnce_peer_reduce_dtext ({(world - 1) * b * K * D * 2 / 1e6:.0f} MB from the peers)               "
                     f"{time_coll(lambda: lib.milnce_peer_reduce_dtext(grads, world, rank, b, K, D, P(dt16), _lib.BF16, st)):8.0f}")
        ring.barrier()
        torch.cuda.synchronize()
    except Exception as e:      # symmetric memory unavailable: keep the rest of the record
        lines.append(f"\n## peer-memory kernels alone: skipped ({type(e).__name__}: {e})")
    if rank == 0:
        text = "\n".join(lines) + "\n"
        print(text)
        if args.out:
            with open(args.out, "w") as f:
                f.write(text)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
