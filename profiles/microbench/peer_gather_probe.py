"""Developer tool (torchrun, >= 2 GPUs): times the pieces of the peer-memory gather path in isolation -
publish kernel, symmetric-memory barrier, staging from the peers' slots (+ the statistics GEMM)."""
import ctypes, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import milnce_b200
from milnce_b200 import _lib
from milnce_b200.utils import PeerRing

world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"]); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, K, D = 8192, 4, 512
b = B // world
v = (torch.randn(b, D) * 0.05).bfloat16().to(dev); t = (torch.randn(b * K, D) * 0.05).bfloat16().to(dev)
lib = _lib.load()
ring = PeerRing(b, K, D, dev)
nbytes = lib.milnce_workspace_bytes(B, K, D, _lib.BF16, b)
ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
pack = torch.empty(2 * b + B, device=dev)
p = lambda x: ctypes.c_void_p(x.data_ptr())
st = _lib.stream_ptr(dev)
arr = (ctypes.c_void_p * world)(*ring.ptrs)

def publish():
    lib.milnce_peer_publish(p(v), p(t), b, K, D, ctypes.c_void_p(ring.ptrs[rank]), ctypes.c_void_p(pack.data_ptr() + 4 * b), st)
def barrier():
    ring.barrier()
def stage_fwd():
    rc = lib.milnce_forward_shard_peers(arr, world, B, K, D, rank * b, b, ctypes.c_void_p(pack.data_ptr() - 4 * rank * b),
                                        ctypes.c_void_p(pack.data_ptr() + 8 * b), p(ws), nbytes, st)
    assert rc == 0
def local_fwd():
    V = torch.empty(B, D, dtype=torch.bfloat16, device=dev); T = torch.empty(B * K, D, dtype=torch.bfloat16, device=dev)
    return V, T
V, T = local_fwd()
def nccl_path_fwd():
    nom = torch.empty(B, device=dev)
    rc = lib.milnce_forward_shard(p(V), p(T), B, K, D, _lib.BF16, rank * b, b, ctypes.c_void_p(pack.data_ptr() - 4 * rank * b),
                                  ctypes.c_void_p(pack.data_ptr() + 8 * b), p(nom), p(ws), nbytes, st)
    assert rc == 0

def timeit(fn, n=30):
    for _ in range(3): fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    x = torch.tensor([e0.elapsed_time(e1) / n * 1e3], device=dev); dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return x.item()

res = [("publish kernel", timeit(publish)), ("symmetric-memory barrier", timeit(barrier)),
       ("stage from peers + statistics GEMM + merges", timeit(stage_fwd)),
       ("local staging + statistics GEMM + merges (milnce_forward_shard on a local gathered copy)", timeit(nccl_path_fwd))]
if rank == 0:
    for k, val in res: print(f"{k:90s} {val:8.0f} us")
dist.destroy_process_group()
