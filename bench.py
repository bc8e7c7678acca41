#!/usr/bin/env python
"""Benchmark of the loss-head hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Headline metric: MIL-NCE forward+backward similarity-pairs/s at global batch
B=8192, K=4, D=512 (BASELINE.json configs[1]; pairs per step = B * B*K).  One
"step" = one forward + one backward of the loss on one batch of synthetic
embeddings.  At N > 1 GPUs the global batch is sharded over ranks (configs[3]),
embeddings are all-gathered with NCCL, each rank computes only its rows of the
similarity matrix and the text gradient is reduce-scattered (strong scaling: total
work fixed).

One JSON line is printed by rank 0.  Besides the contract keys it carries
  roofline      tensor-pipe roofline of the step (algorithmic 6*B*BK*D FLOP)
  cpu_baseline  the oracle's torch port of the reference loss timed on host cores
  sdtw          the soft-DTW half of the metric (DP cells/s, HBM roofline)
--impl reference times the reference's CPU path (oracle port; the reference is
pure Python and /root/reference does not exist on the GPU box).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "milnce_fwd_bwd_similarity_pairs_per_s"
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--candidates", type=int, default=4)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--cpu-sample-batch", type=int, default=2048)
    ap.add_argument("--no-sdtw", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="e2e leg through the eager module instead of the CUDA graph")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"],
                "bf16_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = sorted(float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit())
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_min_mhz": sm[0] if sm else None,
                "sm_max_mhz": mx[0] if mx else None, "power_w": pw[len(pw) // 2] if pw else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


# ----------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle's torch port of loss.py:10-18 on host cores
# ----------------------------------------------------------------------------
def cpu_milnce(batch, K, D, steps, warmup):
    from oracle import milnce_oracle
    torch.manual_seed(0)
    v = (torch.randn(batch, D) * 0.05).requires_grad_(True)
    t = (torch.randn(batch * K, D) * 0.05).requires_grad_(True)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        loss = milnce_oracle.milnce_torch_port(v, t)
        loss.backward()
        v.grad = None
        t.grad = None
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return batch * batch * K / sec, sec


def cpu_sdtw(pairs, N, M, d):
    from oracle import sdtw_oracle
    import numpy as np
    rng = np.random.default_rng(1234)
    x = rng.standard_normal((pairs, N, d)).astype(np.float32)
    y = rng.standard_normal((pairs, M, d)).astype(np.float32)
    sdtw_oracle.softdtw(x[:1], y[:1], gamma=0.1, dist_func="cosine")
    t0 = time.perf_counter()
    sdtw_oracle.softdtw(x, y, gamma=0.1, dist_func="cosine")
    sec = time.perf_counter() - t0
    return 2.0 * pairs * N * M / sec, sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    bs = min(args.cpu_sample_batch, args.batch)
    steps = max(1, min(args.steps, 8))
    warm = max(1, min(args.warmup, 2))
    val, sec = cpu_milnce(bs, args.candidates, args.dim, steps, warm)
    cores = torch.get_num_threads()
    sample = (f"B={bs} of {args.batch}, K={args.candidates}, D={args.dim}, fp32, torch CPU port of "
              f"loss.py:10-18, {steps} steps")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"MILNCELoss fwd+bwd B={args.batch} K={args.candidates} D={args.dim}",
                   "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------
def measured_traffic():
    """DRAM bytes per step from the committed ncu --set full captures (profiles/r1_traffic.json)."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def timed(fn, steps, warmup, barrier):
    for _ in range(warmup):
        fn()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    barrier()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 1e3  # seconds for `steps` steps


def run_b200(args):
    import torch.distributed as dist
    import milnce_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # The JSON line must be the only thing on stdout, and NCCL prints its version banner there:
    # send fd 1 to stderr while the job runs and restore it just before the line is printed.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    B, K, D = args.batch, args.candidates, args.dim
    assert B % world == 0, "global batch must divide by the number of GPUs"
    b = B // world
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    g = torch.Generator().manual_seed(1000 + rank)
    v_host = (torch.randn(b, D, generator=g) * 0.05).to(dtype).pin_memory()
    t_host = (torch.randn(b * K, D, generator=g) * 0.05).to(dtype).pin_memory()
    v_dev = v_host.to(dev).requires_grad_(True)
    t_dev = t_host.to(dev).requires_grad_(True)
    loss_mod = milnce_b200.GatheredMILNCELoss() if world > 1 else milnce_b200.MILNCELoss()

    def step():
        v_dev.grad = None
        t_dev.grad = None
        loss = loss_mod(v_dev, t_dev)
        loss.backward()
        return loss

    # e2e: host buffers in, loss out, through the module API.  Double-buffered like a training input
    # pipeline: step i+1's inputs cross PCIe on a copy stream while step i computes, and every
    # step's H2D copy and D2H loss read happen inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    graphed = world == 1 and not args.no_graph
    if graphed:
        # one CUDA graph per input slot: forward + backward replayed with a single launch
        gsteps = [milnce_b200.GraphedMILNCEStep(B, K, D, dtype=dtype, device=dev) for _ in range(2)]
        bufs = [(gs.video, gs.text) for gs in gsteps]
    else:
        bufs = [(torch.empty_like(v_dev), torch.empty_like(t_dev)) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
    e2e_i = [0]

    def prefetch(slot):
        with torch.cuda.stream(copy_stream), torch.no_grad():
            copy_stream.wait_event(consumed[slot])  # the step that last read this buffer is done
            bufs[slot][0].copy_(v_host, non_blocking=True)
            bufs[slot][1].copy_(t_host, non_blocking=True)
            ready[slot].record(copy_stream)

    for sl in range(2):
        consumed[sl].record(torch.cuda.current_stream())
    prefetch(0)

    def step_e2e():
        slot = e2e_i[0] & 1
        e2e_i[0] += 1
        prefetch(slot ^ 1)
        cur = torch.cuda.current_stream()
        cur.wait_event(ready[slot])
        if graphed:
            loss, _, _ = gsteps[slot].replay()
        else:
            vi = bufs[slot][0].detach().requires_grad_(True)
            ti = bufs[slot][1].detach().requires_grad_(True)
            loss = loss_mod(vi, ti)
            loss.backward()
        consumed[slot].record(cur)
        loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        cur.synchronize()  # the reference's loss.item() (main_distributed.py:241)

    with ClockSampler(local) as clk:
        sec = max_over_ranks(timed(step, args.steps, args.warmup, barrier))
    sec_e2e = max_over_ranks(timed(step_e2e, args.steps, max(3, args.warmup // 2), barrier))

    # forward / backward split (events on the launching stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    fwd_ms = bwd_ms = 0.0
    reps = max(3, args.steps // 4)
    for _ in range(reps):
        v_dev.grad = None
        t_dev.grad = None
        ev[0].record()
        loss = loss_mod(v_dev, t_dev)
        ev[1].record()
        loss.backward()
        ev[2].record()
        torch.cuda.synchronize()
        fwd_ms += ev[0].elapsed_time(ev[1]) / reps
        bwd_ms += ev[1].elapsed_time(ev[2]) / reps

    pk = peaks()
    pairs = float(B) * B * K
    ms_step = sec / args.steps * 1e3
    flop_alg = 6.0 * B * (B * K) * D  # 3 GEMMs, tile recompute not credited
    achieved = flop_alg / (ms_step * 1e-3) / 1e12
    peak = pk["bf16_sustained"] * world
    line = {
        "metric": METRIC, "value": pairs * args.steps / sec, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"MILNCELoss fwd+bwd B={B} K={K} D={D} (BASELINE configs[1])",
                   "global_batch": B, "candidates": K, "dim": D,
                   "parallelism": (f"dp{world}: batch sharded over ranks, NCCL all-gather of the embeddings, each rank "
                                   f"computes its rows of S, reduce-scatter of the text gradient") if world > 1 else "single GPU",
                   "l2": "per-step working set (fp16 G, 0.54 GB written and read twice) exceeds the 126 MB L2; no explicit flush"},
        "e2e": {"value": pairs * args.steps / sec_e2e, "unit": UNIT,
                "h2d_bytes_per_step": (v_host.numel() + t_host.numel()) * v_host.element_size() * world,
                "d2h_bytes_per_step": 4 * world, "ms_per_step": sec_e2e / args.steps * 1e3,
                "pipeline": ("double-buffered inputs: the H2D copy of step i+1 overlaps step i on a copy stream"
                             + ("; forward+backward replayed as one CUDA graph (GraphedMILNCEStep)" if graphed else ""))},
        "gpu_launches": 7 * (args.steps),
        "phases_ms": {"forward": fwd_ms, "backward": bwd_ms},
        "samples_per_s": B * args.steps / sec,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak,
                     "traffic": measured_traffic().get("milnce_step_dram_bytes") if (B, K, D, args.dtype) == (8192, 4, 512, "bf16") else None,
                     "note": (f"algorithmic 6*B*BK*D FLOP per step / step time; peak = bf16 sustained, {pk['src']}; "
                              "traffic = DRAM bytes of one step's four GEMM kernels (ncu --set full, profiles/r1_traffic.json)")},
    }
    if rank == 0:
        line["clocks"] = clk.summary()

    if world == 1 and not args.no_sdtw:
        line["sdtw"] = bench_sdtw(dev, pk, args)
    if rank == 0 and world == 1 and not args.no_cpu:
        bs = min(args.cpu_sample_batch, B)
        val, s = cpu_milnce(bs, K, D, 5, 1)
        line["cpu_baseline"] = {
            "value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"B={bs} of {B}, K={K}, D={D}, fp32, torch CPU port of loss.py:10-18, 5 steps, {s * 1e3:.0f} ms/step"}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        os.dup2(2, 1)
        dist.destroy_process_group()


def bench_sdtw(dev, pk, args):
    """BASELINE configs[2]: 256 pairs of 64x64, d=512, gamma=0.1, cosine, fwd+bwd."""
    import milnce_b200
    Bp, N, M, d = 256, 64, 64, 512
    torch.manual_seed(1234)
    sets = [(torch.randn(Bp, N, d, device=dev).requires_grad_(True),
             torch.randn(Bp, M, d, device=dev).requires_grad_(True)) for _ in range(4)]  # 268 MB > L2
    mod = milnce_b200.SoftDTW(True, gamma=0.1, dist_func="cosine")
    it = [0]

    def eager(x, y):
        x.grad = None
        y.grad = None
        mod(x, y).sum().backward()

    # The step is ~0.16 ms of GPU work behind ~8 launches: from Python it is launch-bound, so each
    # input set's forward+backward is captured once and replayed (same module call, CUDA graph).
    graphs = []
    if not args.no_graph:
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for x, y in sets:
                eager(x, y)
        torch.cuda.current_stream(dev).wait_stream(side)
        for x, y in sets:
            x.grad = None
            y.grad = None
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                mod(x, y).sum().backward()
            graphs.append(g)

    def step():
        k = it[0] % len(sets)
        it[0] += 1
        if graphs:
            graphs[k].replay()
        else:
            eager(*sets[k])

    steps = max(args.steps, 20)
    sec = timed(step, steps, max(args.warmup, 5), lambda: None)
    ms = sec / steps * 1e3
    cells = 2.0 * Bp * N * M
    alg_bytes = 3 * 2 * Bp * N * d * 4 + 2 * Bp * (N + 2) * (M + 2) * 4  # ~210 MB
    out = {"metric": "softdtw_fwd_bwd_dp_cells_per_s", "value": cells / (ms * 1e-3), "unit": "cells/s",
           "ms_per_step": ms, "config": {"workload": "SoftDTW(gamma=0.1, cosine) fwd+bwd 256x(64x64)x512 fp32",
                                         "l2": "4 rotating input sets (268 MB) exceed L2",
                                         "launch": "CUDA graph replay of SoftDTW(...)(x, y).sum().backward()" if graphs else "eager"},
           "roofline": {"bound": "hbm", "achieved": alg_bytes / (ms * 1e-3) / 1e9, "peak": pk["hbm"],
                        "unit": "GB/s", "frac": alg_bytes / (ms * 1e-3) / 1e9 / pk["hbm"],
                        "traffic": measured_traffic().get("sdtw_step_dram_bytes")}}
    if not args.no_cpu:
        val, s = cpu_sdtw(16, N, M, d)
        out["cpu_baseline"] = {"value": val, "unit": "cells/s", "cores": 1, "kind": "port",
                               "sample": f"16 of 256 pairs, oracle (numpy dist + C DP), {s:.2f} s"}
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
