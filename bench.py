#!/usr/bin/env python
"""Benchmark of the loss-head hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Headline metric: MIL-NCE forward+backward similarity-pairs/s at global batch
B=8192, K=4, D=512 (BASELINE.json configs[1]; pairs per step = B * B*K).  One
"step" = one forward + one backward of the loss on one batch of synthetic
embeddings.  At N > 1 GPUs the global batch is sharded over ranks (configs[3]):
every rank obtains the other ranks' text embeddings (read from their peer-mapped
memory on one node, else an NCCL all-gather) and computes its share of the
similarity matrix (strong scaling: total work fixed); rank 0 first checks the
sharded loss and its gradient slices against the single-GPU result.

One JSON line is printed by rank 0.  Besides the contract keys it carries
  roofline            tensor-pipe roofline of the timed region; its peak is the burst or the
                      sustained cuBLAS figure of MEASURED_PEAKS.json, whichever regime the clocks
                      sampled during the region show
  roofline_sustained  the same over a >= 2 s run against the sustained peak (N = 1)
  e2e                 pinned-host inputs in, loss out, through the module API
  cpu_baseline        the reference's own loss.py on the box's host cores (oracle/_ref), else the port
  sdtw                the soft-DTW half of the metric: DP cells/s, HBM roofline, e2e, CPU baseline;
                      at N > 1 every rank runs its own 256 pairs (independent units, weak scaling)
--impl reference times the reference's CPU path on the same config with all host threads.
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
    ap.add_argument("--scheme", default="auto", choices=["auto", "rows", "rowcol"])
    ap.add_argument("--no-sdtw", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time the eager module instead of the CUDA graph")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 2 s sustained-regime leg")
    ap.add_argument("--cpu-budget-s", type=float, default=150.0, help="wall-clock budget of the reference arm")
    ap.add_argument("--config", type=int, default=2, choices=[2, 5],
                    help="2: the loss head alone (the headline, BASELINE configs[1..3]); 5: the S3D-G train step of "
                         "configs[4] with both loss heads (harness/config5.py; needs oracle/_ref)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"],
                "bf16_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "MEASURED_PEAKS.json"}
    return {"hbm": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "src": "B200_PROFILING.md fallback"}


def measure_tf32_peak(dev, seconds=0.4):
    """cuBLAS TF32 GEMM throughput (TFLOP/s), measured here the way MEASURED_PEAKS.json measures bf16: an
    8192^3 torch.matmul with allow_tf32, CUDA events, after warm-up.  Only for --dtype fp32, whose operands are
    float32 (BASELINE.md section 2 asks for the roofline of that mode against a TF32 peak)."""
    n = 8192
    a = torch.randn(n, n, device=dev)
    b = torch.randn(n, n, device=dev)
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        for _ in range(3):
            torch.matmul(a, b)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps, best = 10, 0.0
        elapsed = 0.0
        while elapsed < seconds:
            e0.record()
            for _ in range(reps):
                torch.matmul(a, b)
            e1.record()
            torch.cuda.synchronize(dev)
            t = e0.elapsed_time(e1) / 1e3
            elapsed += t
            best = max(best, 2.0 * n ** 3 * reps / t / 1e12)
        return best
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


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
# reference arm / CPU baselines: the reference's own modules (oracle/_ref) on host cores,
# the oracle's port when they are not on this box
# ----------------------------------------------------------------------------
def host_threads():
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is to use every core it can."""
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0)) or n
    except Exception:
        pass
    torch.set_num_threads(n)
    return n


def cpu_milnce(batch, K, D, steps, warmup, budget_s):
    """fp32 MILNCELoss forward+backward on the host.  Returns (pairs/s, s/step, steps, warmup, kind)."""
    from oracle import milnce_oracle, ref_loader
    kind = "port"
    fn = milnce_oracle.milnce_torch_port
    shim = None
    if not ref_loader.why_unavailable():
        mod = ref_loader.load("loss").MILNCELoss()   # loss.py:6-18, unmodified
        fn = lambda v, t: mod(v, t)
        shim = ref_loader.cpu_cuda_shim
        kind = "reference"
    torch.manual_seed(0)
    v = (torch.randn(batch, D) * 0.05).requires_grad_(True)
    t = (torch.randn(batch * K, D) * 0.05).requires_grad_(True)

    def one():
        t0 = time.perf_counter()
        if shim is not None:
            with shim():
                loss = fn(v, t)
        else:
            loss = fn(v, t)
        loss.backward()
        v.grad = None
        t.grad = None
        return time.perf_counter() - t0

    t_start = time.perf_counter()
    first = one()
    warm_done = 1
    while warm_done < warmup and (time.perf_counter() - t_start) + first * 2 < budget_s * 0.25:
        one()
        warm_done += 1
    times = []
    while len(times) < steps:
        times.append(one())
        if (time.perf_counter() - t_start) + times[-1] > budget_s:
            break
    sec = sum(times) / len(times)
    return batch * batch * K / sec, sec, len(times), warm_done, kind


def cpu_sdtw(pairs, N, M, d):
    """SoftDTW(gamma=0.1, cosine) forward+backward on the host.  Returns (cells/s, seconds, kind, note)."""
    from oracle import ref_loader
    if not ref_loader.why_unavailable():
        sd = ref_loader.load("soft_dtw_cuda")
        mod = sd.SoftDTW(False, 0.1, dist_func="cosine")     # numba-CPU path, soft_dtw_cuda.py:243-271
        torch.manual_seed(1234)
        x = torch.randn(pairs, N, d, requires_grad=True)
        y = torch.randn(pairs, M, d, requires_grad=True)
        mod(x[:2], y[:2]).sum().backward()                   # numba JIT outside the timed region
        x.grad = y.grad = None
        t0 = time.perf_counter()
        mod(x, y).sum().backward()
        sec = time.perf_counter() - t0
        return 2.0 * pairs * N * M / sec, sec, "reference", "soft_dtw_cuda.py numba-CPU path with its torch distance function"
    from oracle import sdtw_oracle
    import numpy as np
    rng = np.random.default_rng(1234)
    x = rng.standard_normal((pairs, N, d)).astype(np.float32)
    y = rng.standard_normal((pairs, M, d)).astype(np.float32)
    sdtw_oracle.softdtw(x[:1], y[:1], gamma=0.1, dist_func="cosine")
    t0 = time.perf_counter()
    sdtw_oracle.softdtw(x, y, gamma=0.1, dist_func="cosine")
    sec = time.perf_counter() - t0
    return 2.0 * pairs * N * M / sec, sec, "port", "oracle (numpy distance + C recurrences)"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_threads()
    B, K, D = args.batch, args.candidates, args.dim
    val, sec, steps, warm, kind = cpu_milnce(B, K, D, args.steps, args.warmup, args.cpu_budget_s)
    what = "reference loss.py:6-18 (unmodified, .cuda() shimmed to the host)" if kind == "reference" else \
        "torch port of loss.py:10-18 (oracle/_ref not on this box)"
    sample = (f"B={B}, K={K}, D={D}, fp32, {what}, {steps} steps after {warm} warm-up "
              f"(requested {args.steps}/{args.warmup}; capped by a {args.cpu_budget_s:.0f} s budget)")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"MILNCELoss fwd+bwd B={B} K={K} D={D} (BASELINE configs[1])",
                   "global_batch": B, "candidates": K, "dim": D, "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------
def measured_traffic():
    """DRAM bytes per step from the committed ncu --set full captures (newest round first)."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                d = json.load(f)
                d["_file"] = "profiles/" + name
                return d
        except Exception:
            continue
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


def launches():
    import milnce_b200
    return int(milnce_b200._lib.load().b200head_launch_count())


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

    def shard(r):
        g = torch.Generator().manual_seed(1000 + r)
        return ((torch.randn(b, D, generator=g) * 0.05).to(dtype), (torch.randn(b * K, D, generator=g) * 0.05).to(dtype))

    v_host, t_host = (x.pin_memory() for x in shard(rank))
    v_dev = v_host.to(dev).requires_grad_(True)
    t_dev = t_host.to(dev).requires_grad_(True)
    loss_mod = milnce_b200.GatheredMILNCELoss(scheme=args.scheme) if world > 1 else milnce_b200.MILNCELoss()

    def step():
        v_dev.grad = None
        t_dev.grad = None
        loss = loss_mod(v_dev, t_dev)
        loss.backward()
        return loss

    # ---- N > 1: the sharded result against the single-GPU module, before anything is timed ----
    parity = None
    if world > 1:
        loss_s = step().detach().clone()
        dv_s, dt_s = v_dev.grad.clone(), t_dev.grad.clone()
        ok = torch.ones(1, device=dev)
        if rank == 0:
            full = [shard(r) for r in range(world)]
            V = torch.cat([f[0] for f in full]).to(dev)
            T = torch.cat([f[1] for f in full]).to(dev)
            # single-GPU result with FLOAT32 gradient outputs straight from the C ABI (the sharded module's bf16
            # gradients are then one bf16 rounding, 1.7e-3 norm-wise, plus the rounding of the partials away)
            from milnce_b200.loss import CudaBackend
            Vd, Td = V.detach(), T.detach()
            ref_l, ref_den, ref_nom, ref_ws = CudaBackend.forward(Vd, Td, 0, B)
            ref_dv, ref_dt = CudaBackend.backward(Vd, Td, ref_den, ref_nom, torch.ones(1, device=dev), 0, B, ref_ws)
            ref = ref_l[0]
            rel = lambda a, r_: ((a.double() - r_.double()).norm() / r_.double().norm()).item()
            e = {"loss": abs(loss_s.item() - ref.item()) / abs(ref.item()),
                 "dvideo": rel(dv_s, ref_dv[:b]), "dtext": rel(dt_s, ref_dt[:b * K])}
            del ref_dv, ref_dt, ref_ws
            gtol = 3e-3 if dtype == torch.bfloat16 else 1e-5
            passed = e["loss"] < 1e-5 and e["dvideo"] < gtol and e["dtext"] < gtol
            parity = {"parity_checked": True, "passed": bool(passed), "errors": e,
                      "tolerance": {"loss": 1e-5, "grad": gtol},
                      "what": "rank 0: sharded loss and its own gradient slices vs the single-GPU C-ABI result on the whole batch (float32 gradient outputs)"}
            del V, T, ref, full, Vd, Td
            torch.cuda.empty_cache()
            if not passed:
                ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() == 0:
            if rank == 0:
                print(f"bench.py: multi-GPU parity check FAILED: {parity}", file=sys.stderr)
            sys.exit(3)

    # ---- device-resident leg (`value`) ----
    # forward + backward replayed as one CUDA graph through the package's GraphedMILNCEStep (N = 1) or
    # GraphedGatheredMILNCEStep (N > 1: the NCCL collectives are part of the graph); the eager module is
    # also timed and reported.
    graphed = not args.no_graph and (world == 1 or dtype == torch.bfloat16)
    gsteps = []
    launches_per_replay = None
    if graphed:
        n0 = launches()
        if world == 1:
            gsteps = [milnce_b200.GraphedMILNCEStep(B, K, D, dtype=dtype, device=dev, warmup=1) for _ in range(2)]
        else:
            gsteps = [milnce_b200.GraphedGatheredMILNCEStep(b, K, D, dtype=dtype, device=dev, scheme=args.scheme,
                                                            warmup=1) for _ in range(2)]
        launches_per_replay = (launches() - n0) // (2 * 2)   # per instance: one warm-up pass + the capture
        with torch.no_grad():
            for gs in gsteps:
                gs.video.copy_(v_dev)
                gs.text.copy_(t_dev)

    def step_graph():
        gsteps[0].replay()

    n0 = launches()
    with ClockSampler(local) as clk_e:
        sec_eager = max_over_ranks(timed(step, args.steps, args.warmup, barrier))
    eager_launches = launches() - n0 - 0
    eager_launches = eager_launches * args.steps // (args.steps + args.warmup)
    clocks = clk_e.summary()
    sec = sec_eager
    gpu_launches = eager_launches
    if graphed:
        with ClockSampler(local) as clk_g:
            sec = timed(step_graph, args.steps, args.warmup, barrier)
        clocks = clk_g.summary()
        gpu_launches = launches_per_replay * args.steps

    # ---- sustained regime: >= 2 s of back-to-back steps (the GPU settles at its power cap) ----
    sustained = None
    if world == 1 and not args.no_sustained:
        per = sec / args.steps
        n_sus = max(args.steps, int(2.2 / per))
        with ClockSampler(local) as clk_s:
            sec_s = timed(step_graph if graphed else step, n_sus, 3, barrier)
        sustained = (sec_s / n_sus, n_sus, clk_s.summary())

    # ---- e2e: host buffers in, loss out, through the module API.  Double-buffered like a training
    # input pipeline: step i+1's inputs cross PCIe on a copy stream while step i computes, and
    # every step's H2D copy and D2H loss read happen inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    if graphed:
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

    sec_e2e = max_over_ranks(timed(step_e2e, args.steps, max(3, args.warmup // 2), barrier))

    # forward / backward split of the eager module (events on the launching stream)
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
    # the peak of the regime the timed region ran in: a short region without a power cap is a burst
    capped = "sw_power_cap" in (clocks.get("reasons") or [])
    burst = (sec < 1.0) and not capped
    peak1 = pk["bf16_burst"] if burst else pk["bf16_sustained"]
    peak = peak1 * world
    traffic = measured_traffic()
    line = {
        "metric": METRIC, "value": pairs * args.steps / sec, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"MILNCELoss fwd+bwd B={B} K={K} D={D} (BASELINE configs[1])",
                   "global_batch": B, "candidates": K, "dim": D,
                   "launch": (("one CUDA graph replay per step (" + type(gsteps[0]).__name__ + ")") if graphed
                              else "eager module calls"),
                   "parallelism": (f"dp{world}: batch sharded over ranks, embeddings via {getattr(loss_mod, 'last_gather', None)} "
                                   f"(peer = read from the peers' symmetric memory, nccl = one grouped all-gather), "
                                   f"scheme={getattr(loss_mod, 'last_scheme', args.scheme)}"
                                   if world > 1 else "single GPU"),
                   "l2": "per-step working set (fp16 G, 0.54 GB written and read twice) exceeds the 126 MB L2; no explicit flush"},
        "e2e": {"value": pairs * args.steps / sec_e2e, "unit": UNIT,
                "h2d_bytes_per_step": (v_host.numel() + t_host.numel()) * v_host.element_size() * world,
                "d2h_bytes_per_step": 4 * world, "ms_per_step": sec_e2e / args.steps * 1e3,
                "pipeline": ("double-buffered inputs: the H2D copy of step i+1 overlaps step i on a copy stream"
                             + ("; forward+backward replayed as one CUDA graph" if graphed else ""))},
        "gpu_launches": int(gpu_launches),
        "gpu_launches_how": ("b200head_launch_count() over the graph capture x replays" if graphed
                             else "b200head_launch_count() difference over warm-up + timed steps, scaled to the timed steps"),
        "eager": {"ms_per_step": sec_eager / args.steps * 1e3, "gpu_launches": int(eager_launches)},
        "phases_ms": {"forward": fwd_ms, "backward": bwd_ms},
        "samples_per_s": B * args.steps / sec,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak,
                     "regime": "burst" if burst else "sustained",
                     "traffic": traffic.get("milnce_step_dram_bytes") if (B, K, D, args.dtype) == (8192, 4, 512, "bf16") else None,
                     "note": (f"algorithmic 6*B*BK*D FLOP per step / step time; peak = cuBLAS bf16 "
                              f"{'burst' if burst else 'sustained'} figure of {pk['src']} (timed region {sec:.3f} s, "
                              f"power cap {'seen' if capped else 'not seen'}); traffic = DRAM bytes of one step "
                              f"(ncu --set full, {traffic.get('_file')})")},
    }
    if args.dtype == "fp32":
        tf32 = measure_tf32_peak(dev) if rank == 0 or world > 1 else 0.0
        line["roofline"].update({"peak": tf32 * world, "frac": achieved / (tf32 * world),
                                 "note": "float32 inputs: algorithmic 6*B*BK*D FLOP per step against the cuBLAS TF32 "
                                         "peak measured in this process (8192^3 matmul, allow_tf32, best of >= 0.4 s); "
                                         "the path itself runs 3 fp16 planes on the kind::f16 tensor pipe"})
    if sustained is not None and args.dtype != "fp32":   # (the sustained peak on file is a bf16 figure)
        per_s, n_s, clk_sum = sustained
        ach_s = flop_alg / per_s / 1e12
        line["roofline_sustained"] = {"bound": "tensor", "achieved": ach_s, "peak": pk["bf16_sustained"],
                                      "unit": "TFLOP/s", "frac": ach_s / pk["bf16_sustained"],
                                      "ms_per_step": per_s * 1e3, "steps": n_s, "clocks": clk_sum}
    if parity is not None:
        line.update({"parity_checked": True, "parity": parity})
    if rank == 0:
        line["clocks"] = clocks

    if not args.no_sdtw:
        sd = bench_sdtw(dev, pk, args, world, rank, barrier, max_over_ranks)
        if rank == 0:
            line["sdtw"] = sd
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = host_threads()
        val, s, n, w, kind = cpu_milnce(B, K, D, 2, 1, 30.0)
        line["cpu_baseline"] = {
            "value": val, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": (f"B={B}, K={K}, D={D}, fp32, "
                       + ("reference loss.py (oracle/_ref)" if kind == "reference" else "torch port of loss.py:10-18")
                       + f", {n} steps after {w} warm-up, {s * 1e3:.0f} ms/step")}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        os.dup2(2, 1)
        for gs in gsteps:
            gs.close()
        dist.destroy_process_group()


def bench_sdtw(dev, pk, args, world, rank, barrier, max_over_ranks):
    """BASELINE configs[2]: 256 pairs of 64x64, d=512, gamma=0.1, cosine, fwd+bwd, on EVERY rank
    (pairs are independent units: no collective; value = pairs of all ranks / max-over-ranks time)."""
    import milnce_b200
    Bp, N, M, d = 256, 64, 64, 512
    torch.manual_seed(1234 + rank)
    sets = [(torch.randn(Bp, N, d, device=dev).requires_grad_(True),
             torch.randn(Bp, M, d, device=dev).requires_grad_(True)) for _ in range(4)]  # 268 MB > L2
    mod = milnce_b200.SoftDTW(True, gamma=0.1, dist_func="cosine")
    it = [0]

    def eager(x, y):
        x.grad = None
        y.grad = None
        mod(x, y).sum().backward()

    # The step is well under 0.2 ms of GPU work behind several launches: from Python it is launch-bound,
    # so each input set's forward+backward is captured once and replayed (same module call, CUDA graph).
    graphs = []
    per_replay = 0
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
            n0 = launches()
            with torch.cuda.graph(g):
                mod(x, y).sum().backward()
            per_replay = launches() - n0
            graphs.append(g)

    def step():
        k = it[0] % len(sets)
        it[0] += 1
        if graphs:
            graphs[k].replay()
        else:
            eager(*sets[k])

    steps = max(args.steps, 20)
    sec = max_over_ranks(timed(step, steps, max(args.warmup, 5), barrier))
    ms = sec / steps * 1e3
    cells = 2.0 * Bp * N * M * world
    alg_bytes = (3 * 2 * Bp * N * d * 4 + 2 * Bp * (N + 2) * (M + 2) * 4) * world  # ~210 MB per GPU

    # e2e: pinned host X, Y in; out, dX, dY back to pinned host; every step, inside the timed region
    xh = torch.randn(Bp, N, d).pin_memory()
    yh = torch.randn(Bp, M, d).pin_memory()
    oh = torch.empty(Bp).pin_memory()
    dxh = torch.empty(Bp, N, d).pin_memory()
    dyh = torch.empty(Bp, M, d).pin_memory()
    xd, yd = sets[0]
    kgraph = graphs[0] if graphs else None
    out_static = [None]
    if kgraph is not None:
        # a graph whose output tensor is kept, for the D2H read
        xd.grad = None
        yd.grad = None
        kgraph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(kgraph):
            out_static[0] = mod(xd, yd)
            out_static[0].sum().backward()

    def step_e2e():
        with torch.no_grad():
            xd.copy_(xh, non_blocking=True)
            yd.copy_(yh, non_blocking=True)
        if kgraph is not None:
            kgraph.replay()
            o = out_static[0]
        else:
            xd.grad = None
            yd.grad = None
            o = mod(xd, yd)
            o.sum().backward()
        oh.copy_(o.detach(), non_blocking=True)
        dxh.copy_(xd.grad, non_blocking=True)
        dyh.copy_(yd.grad, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    sec_e = max_over_ranks(timed(step_e2e, steps, 3, barrier))
    h2d = 2 * Bp * N * d * 4
    d2h = 2 * Bp * N * d * 4 + Bp * 4
    traffic = measured_traffic()
    out = {"metric": "softdtw_fwd_bwd_dp_cells_per_s", "value": cells / (ms * 1e-3), "unit": "cells/s",
           "ms_per_step": ms, "n_gpus": world, "scaling": "weak",
           "config": {"workload": "SoftDTW(gamma=0.1, cosine) fwd+bwd 256x(64x64)x512 fp32 per GPU (BASELINE configs[2])",
                      "l2": "4 rotating input sets (268 MB) exceed L2",
                      "launch": "CUDA graph replay of SoftDTW(...)(x, y).sum().backward()" if graphs else "eager",
                      "parallelism": f"{world} x 256 independent pairs, no collective" if world > 1 else "single GPU"},
           "gpu_launches": int(per_replay * steps) if graphs else None,
           "e2e": {"value": cells / (sec_e / steps), "unit": "cells/s", "ms_per_step": sec_e / steps * 1e3,
                   "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                   "pipeline": "pinned host X, Y -> device, fwd+bwd, out and dX, dY -> pinned host; not overlapped"},
           "roofline": {"bound": "hbm", "achieved": alg_bytes / (ms * 1e-3) / 1e9, "peak": pk["hbm"] * world,
                        "unit": "GB/s", "frac": alg_bytes / (ms * 1e-3) / 1e9 / (pk["hbm"] * world),
                        "traffic": traffic.get("sdtw_step_dram_bytes")}}
    if world == 1:
        # the general path (soft_dtw_cuda.py's own use: euclidean, longer sequences): eager, events, fwd+bwd
        gen = []
        for (bp, n, m, dd) in ((64, 256, 256, 64), (8, 1024, 1024, 64)):
            gx = torch.randn(bp, n, dd, device=dev).requires_grad_(True)
            gy = torch.randn(bp, m, dd, device=dev).requires_grad_(True)
            gmod = milnce_b200.SoftDTW(True, gamma=1.0, dist_func="euclidean")

            def gstep():
                gx.grad = None
                gy.grad = None
                gmod(gx, gy).sum().backward()
            gsec = timed(gstep, 10, 3, barrier) / 10
            gen.append({"pairs": bp, "N": n, "M": m, "d": dd, "dist_func": "euclidean", "gamma": 1.0,
                        "ms_per_step": gsec * 1e3, "cells_per_s": 2.0 * bp * n * m / gsec})
        out["general"] = gen
    if rank == 0 and world == 1 and not args.no_cpu:
        val, s, kind, note = cpu_sdtw(Bp, N, M, d)
        out["cpu_baseline"] = {"value": val, "unit": "cells/s", "cores": host_threads(), "kind": kind,
                               "sample": f"all {Bp} pairs once, {note}, {s:.2f} s (the recurrences are single-threaded)"}
    return out


def main():
    args = parse()
    if args.config == 5:
        sys.argv = [os.path.join(ROOT, "harness", "config5.py"), "--steps", str(max(2, min(args.steps, 6))),
                    "--warmup", str(max(1, min(args.warmup, 2)))]
        import runpy
        runpy.run_path(sys.argv[0], run_name="__main__")
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
