"""Host time of one eager MILNCELoss forward+backward (no synchronisation inside the loop) against
the GPU time of the same steps: tells whether the eager module is launch/host-bound."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import milnce_b200  # noqa: E402

B, K, D = 8192, 4, 512
v = (torch.randn(B, D) * 0.05).bfloat16().cuda().requires_grad_(True)
t = (torch.randn(B * K, D) * 0.05).bfloat16().cuda().requires_grad_(True)
mod = milnce_b200.MILNCELoss()
for _ in range(5):
    v.grad = t.grad = None
    mod(v, t).backward()
torch.cuda.synchronize()
n = 50
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(n):
    v.grad = t.grad = None
    mod(v, t).backward()
e1.record()
host = (time.perf_counter() - t0) / n
torch.cuda.synchronize()
print(f"eager step: host issue time {host * 1e6:.0f} us, GPU time {e0.elapsed_time(e1) / n * 1e3:.0f} us")
# forward only / backward only host cost
t0 = time.perf_counter()
for _ in range(n):
    loss = mod(v, t)
host_f = (time.perf_counter() - t0) / n
torch.cuda.synchronize()
print(f"forward only: host {host_f * 1e6:.0f} us")
step = milnce_b200.GraphedMILNCEStep(B, K, D)
torch.cuda.synchronize()
e0.record()
for _ in range(n):
    step.replay()
e1.record()
torch.cuda.synchronize()
print(f"graph replay: GPU time {e0.elapsed_time(e1) / n * 1e3:.0f} us")
