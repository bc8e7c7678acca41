"""soft-DTW forward+backward at the BASELINE shape, timed like bench.py's soft-DTW leg (4 rotating input sets,
268 MB > L2, one CUDA graph per set, CUDA events).  Loads the library B200HEAD_LIB names, for A/B of variants."""
import sys, torch
sys.path.insert(0, "/root/repo")
import milnce_b200
torch.manual_seed(1234)
sets = [(torch.randn(256, 64, 512, device="cuda", requires_grad=True),
         torch.randn(256, 64, 512, device="cuda", requires_grad=True)) for _ in range(4)]
mod = milnce_b200.SoftDTW(True, gamma=0.1, dist_func="cosine")
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for x, y in sets:
        mod(x, y).sum().backward()
torch.cuda.current_stream().wait_stream(side)
graphs = []
for x, y in sets:
    x.grad = None; y.grad = None
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        mod(x, y).sum().backward()
    graphs.append(g)
res = []
for rep in range(5):
    for i in range(8):
        graphs[i % 4].replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40):
        graphs[i % 4].replay()
    e1.record()
    torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) / 40)
print("ms per fwd+bwd, 5 runs of 40 replays:", " ".join(f"{r:.4f}" for r in res))
