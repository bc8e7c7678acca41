"""Is the gradient GEMM bound by DRAM reads of A (128-byte pieces at a 64 KB row pitch)?
Times the wide kernel on the dvideo shape, and on a shape whose A fits in L2 (second run: A comes from L2)."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import milnce_b200  # noqa: E402

lib = milnce_b200._lib.load()
scratch = torch.empty(lib.b200head_gemm_wide_scratch_bytes(), dtype=torch.uint8, device="cuda")


def run(M, N, Kd, transposed, reps=10, wide=True):
    b = torch.randn(N, Kd, device="cuda").half()
    a = (torch.randn(Kd, M, device="cuda") if transposed else torch.randn(M, Kd, device="cuda")).half()
    c = torch.empty(M, N, dtype=torch.float32, device="cuda")
    st = milnce_b200._lib.stream_ptr()

    def call():
        if wide:
            rc = lib.b200head_gemm_wide_f16(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                            ctypes.c_void_p(c.data_ptr()), M, N, Kd, a.stride(0), b.stride(0), c.stride(0),
                                            ctypes.c_float(1.0), transposed, ctypes.c_void_p(scratch.data_ptr()), st)
        else:
            rc = lib.b200head_gemm2_f16(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                        ctypes.c_void_p(c.data_ptr()), M, N, Kd, a.stride(0), b.stride(0), c.stride(0),
                                        ctypes.c_float(1.0), transposed, st)
        assert rc == 0
    for _ in range(3):
        call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        call()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    tf = 2.0 * M * N * Kd / us / 1e6
    print(f"{'wide' if wide else 'pair'} M={M:6d} N={N} Kd={Kd:6d} t={transposed}: {us:8.1f} us  {tf:7.1f} TFLOP/s  A={M * Kd * 2 / 1e6:.0f} MB")


# equal FLOPs and equal bytes of A (537 MB, streamed from DRAM); only the row pitch of A (= 2*Kd bytes) and
# the tile count change: 64 / 128 / 256 / 512 tiles for 74 CTA pairs
for M, Kd in ((8192, 32768), (16384, 16384), (32768, 8192), (65536, 4096), (131072, 2048)):
    run(M, 512, Kd, 0, wide=False, reps=20)
