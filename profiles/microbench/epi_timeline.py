"""Timeline of CTA 0's warps in the CTA-pair GEMM kernels (instrumented build, -DB200_EPI_STAMP).
usage: B200HEAD_LIB=profiles/microbench/libb200head_stamp.so python profiles/microbench/epi_timeline.py [-v]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import milnce_b200  # noqa: E402

KINDS, WARPS, TILES, SLOTS = 3, 20, 64, 16
lib = ctypes.CDLL(os.environ["B200HEAD_LIB"])
lib.b200head_debug_stamps.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
VERBOSE = "-v" in sys.argv


def stamps():
    a = np.zeros(KINDS * WARPS * TILES * SLOTS, dtype=np.int64)
    assert lib.b200head_debug_stamps(a.ctypes.data, a.size) == 0
    return a.reshape(KINDS, WARPS, TILES, SLOTS)


def summary(name, s, nwarps, lo=10, hi=50):
    """s: (WARPS, TILES, SLOTS).  Warp 1 = MMA issuer (slots 0,1,2 = tile start, tempty arrived, commit);
    warps 2.. = epilogue (0 start, 1 pre-wait, 2 tfull seen, 3 ld0, 4 run0 end, 5 ld1, 6 run1 end)."""
    mma = s[1]
    period = np.diff(mma[lo:hi + 1, 0]).mean()
    wait_tempty = (mma[lo:hi, 1] - mma[lo:hi, 0]).mean()
    mma_busy = (mma[lo:hi, 2] - mma[lo:hi, 1]).mean()
    e = s[2:2 + nwarps, lo:hi, :]
    epi_wait = (e[:, :, 2] - e[:, :, 1]).mean()
    epi_busy = (e[:, :, 6] - e[:, :, 2]).mean()
    run0 = (e[:, :, 4] - e[:, :, 3]).mean()
    run1 = (e[:, :, 6] - e[:, :, 5]).mean()
    ld = ((e[:, :, 3] - e[:, :, 2]) + (e[:, :, 5] - e[:, :, 4])).mean()
    print(f"{name:28s} tile period {period:7.0f} clk | MMA: wait tempty {wait_tempty:6.0f}, issue->commit {mma_busy:6.0f} | "
          f"epilogue: wait tfull {epi_wait:6.0f}, busy {epi_busy:6.0f} (ld {ld:5.0f}, run0 {run0:5.0f}, run1 {run1:5.0f})")


def show(name, s, nwarps, slots, tiles=range(20, 23)):
    print(f"== {name}: cycles relative to the MMA warp's stamp 0 of tile {tiles[0]}")
    base = s[1, tiles[0], 0]
    for t in tiles:
        mma = s[1, t, :3] - base
        print(f"tile {t}: MMA wait-tempty {mma[0]}..{mma[1]} commit {mma[2]}")
        for w in range(2, 2 + nwarps, max(1, nwarps // 4)):
            e = s[w, t, :] - base
            print(f"    warp {w:2d}: " + " ".join(f"{n}={e[i]}" for i, n in slots))


B, K, D = 8192, 4, 512
torch.manual_seed(0)
v = (torch.randn(B, D, device="cuda") * 0.05).bfloat16().requires_grad_(True)
t = (torch.randn(B * K, D, device="cuda") * 0.05).bfloat16().requires_grad_(True)
mod = milnce_b200.MILNCELoss()
for _ in range(3):
    loss = mod(v, t)
    torch.cuda.synchronize()
    s_fwd = stamps()
    loss.backward()
    torch.cuda.synchronize()
    s_bwd = stamps()
nrc = int(os.environ.get("ROWCOL_WARPS", "8"))
print(os.path.basename(os.environ["B200HEAD_LIB"]))
summary("EpiRowCol", s_fwd[1], nrc)
summary("EpiG", s_bwd[2], 16)
if VERBOSE:
    base_slots = [(0, "start"), (1, "pre-wait"), (2, "tfull"), (3, "ld0"), (4, "run0"), (5, "ld1"), (6, "run1")]
    show("EpiRowCol", s_fwd[1], nrc, base_slots)
    show("EpiG", s_bwd[2], 16, base_slots + [(8, "tab0"), (9, "math0"), (10, "tab1"), (11, "math1")])
    show("EpiStore dT", s_bwd[0], 8, base_slots, tiles=range(1, 3))
