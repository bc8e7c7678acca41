"""soft-DTW forward+backward on 4 rotating input sets (268 MB > L2), like bench.py's soft-DTW leg: for ncu."""
import sys, torch
sys.path.insert(0, "/root/repo")
import milnce_b200
torch.manual_seed(1234)
sets = [(torch.randn(256, 64, 512, device="cuda", requires_grad=True),
         torch.randn(256, 64, 512, device="cuda", requires_grad=True)) for _ in range(4)]
mod = milnce_b200.SoftDTW(True, gamma=0.1, dist_func="cosine")
for it in range(12):
    x, y = sets[it % 4]
    x.grad = None; y.grad = None
    mod(x, y).sum().backward()
torch.cuda.synchronize()
print("ok")
