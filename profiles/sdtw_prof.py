import sys, torch
sys.path.insert(0, "/root/repo")
import milnce_b200
torch.manual_seed(1234)
x = torch.randn(256, 64, 512, device="cuda", requires_grad=True)
y = torch.randn(256, 64, 512, device="cuda", requires_grad=True)
mod = milnce_b200.SoftDTW(True, gamma=0.1, dist_func="cosine")
for _ in range(6):
    x.grad = None; y.grad = None
    mod(x, y).sum().backward()
torch.cuda.synchronize()
print("ok")
