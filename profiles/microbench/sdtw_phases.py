"""Developer tool: phase timestamps of the fused soft-DTW kernels (profiling build of the library)."""
import ctypes, os, sys
import torch
here = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.environ.get("B200HEAD_LIB") or os.path.join(here, "libb200head_prof.so"))
lib.sdtw_aux_bytes.restype = ctypes.c_size_t
B, N, M, d = int(os.environ.get('SDTW_PAIRS', '256')), 64, 64, 512
torch.manual_seed(0)
x = torch.randn(B, N, d, device="cuda"); y = torch.randn(B, M, d, device="cuda")
nb = lib.sdtw_aux_bytes(B, N, M, d)
aux = torch.empty(nb, dtype=torch.uint8, device="cuda")
out = torch.empty(B, device="cuda"); go = torch.ones(B, device="cuda")
dx = torch.empty_like(x); dy = torch.empty_like(y)
p = lambda t: ctypes.c_void_p(t.data_ptr())
for _ in range(3):
    rc = lib.sdtw_forward(p(x), p(y), B, N, M, d, ctypes.c_float(0.1), ctypes.c_float(0.0), 0, None, p(out), p(aux), ctypes.c_size_t(nb), None)
    assert rc == 0
    rc = lib.sdtw_backward(p(x), p(y), p(go), B, N, M, d, ctypes.c_float(0.1), ctypes.c_float(0.0), 0, p(aux), ctypes.c_size_t(nb), None, p(dx), p(dy), None)
    assert rc == 0
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 32)()
lib.sdtw_debug_read(buf)
t = list(buf)
print("fwd: GEMM phase %d  (split thread waited for TMA %d, drain thread waited for MMAs %d)  join %d  tile %d  dp %d  weights+R %d  (cycles, block 0)"
      % (t[1] - t[0], t[5], t[6], t[2] - t[1], t[7] - t[2], t[3] - t[7], t[4] - t[3]))
print("pairs", B)
print("fwd wavefront: leader warp done after %d, follower after %d; bwd wavefront: leader (warp 1) after %d, follower after %d" % (t[12] - t[7], t[3] - t[7], t[13] - t[8], t[9] - t[8]))
print("bwd: dp %d  scale %d  gemm %d" % (t[9] - t[8], t[10] - t[9], t[11] - t[10]))
print("bwd GEMM phase: drain thread: waiting for MMAs %d, TMEM loads %d, barriers %d, exchange+global %d; split thread waited for TMA %d; TMA warp waited for free slots %d" % (t[16], t[17], t[18], t[19], t[20], t[21]))
z = t[22]
print("bwd item 5 timeline (cycles after its TMA issue): split sees data %d, split work done %d, split arrived %d | MMA warp: accumulator free %d, operands ready %d, MMAs issued+committed %d, MMAs COMPLETE %d | drain: sees accumulator %d, TMEM read %d, item done %d"
      % (t[23] - z, t[30] - z, t[24] - z, t[31] - z, t[25] - z, t[26] - z, t[14] - z, t[27] - z, t[28] - z, t[29] - z))
