"""Developer tool: phase timestamps of the fused soft-DTW kernels (profiling build of the library)."""
import ctypes, os, sys
import torch
here = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.environ.get("B200HEAD_LIB") or os.path.join(here, "libb200head_prof.so"))
lib.sdtw_aux_bytes.restype = ctypes.c_size_t
B, N, M, d = 256, 64, 64, 512
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
buf = (ctypes.c_longlong * 16)()
lib.sdtw_debug_read(buf)
t = list(buf)
print("fwd: mma %d  tile %d  dp %d  weights+R %d  (cycles, block 0)" % (t[1] - t[0], t[2] - t[1], t[3] - t[2], t[4] - t[3]))
print("bwd: dp %d  scale %d  gemm %d" % (t[9] - t[8], t[10] - t[9], t[11] - t[10]))
