"""The peer-memory entry points (include/b200head.h: milnce_peer_publish, milnce_forward_shard_peers,
milnce_peer_reduce_dtext) driven on ONE GPU: the W "ranks" are W slots / partial buffers in the same device
memory, so the kernels run exactly as on W GPUs except that the loads do not cross NVLink.  (The real
two-process run over symmetric memory is tests/test_gpu_multi.py.)"""
import ctypes

import numpy as np
import pytest
import torch

from milnce_b200 import _lib
from milnce_b200.loss import CudaBackend
from oracle import milnce_oracle

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


@pytest.mark.parametrize("W,b,K,D", [(2, 64, 4, 512), (4, 64, 2, 72), (3, 128, 1, 256), (4, 128, 2, 64), (3, 256, 4, 128)])
def test_peer_forward_and_reduce_match_the_gathered_path(W, b, K, D):
    lib = _lib.load()
    B = W * b
    g = torch.Generator().manual_seed(W * 1000 + b + D)
    V = (torch.randn(B, D, generator=g) * 0.2).bfloat16().cuda()
    T = (torch.randn(B * K, D, generator=g) * 0.2).bfloat16().cuda()
    st = _lib.stream_ptr()
    slot_bytes = lib.milnce_peer_slot_bytes(b, K, D)
    slots = [torch.zeros(slot_bytes, dtype=torch.uint8, device="cuda") for _ in range(W)]
    noms = torch.empty(B, device="cuda")
    for r in range(W):
        rc = lib.milnce_peer_publish(_p(V[r * b:(r + 1) * b]), _p(T[r * b * K:(r + 1) * b * K]), b, K, D, _p(slots[r]),
                                     ctypes.c_void_p(noms.data_ptr() + 4 * r * b), st)
        _lib.check(rc, "milnce_peer_publish")
    arr = (ctypes.c_void_p * W)(*[s.data_ptr() for s in slots])
    packs, wss = [], []
    nbytes = lib.milnce_workspace_bytes(B, K, D, _lib.BF16, b)
    for r in range(W):
        ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        row = torch.full((B,), float("nan"), device="cuda")
        col = torch.empty(B, device="cuda")
        rc = lib.milnce_forward_shard_peers(arr, W, B, K, D, r * b, b, _p(row), _p(col), _p(ws), nbytes, st)
        _lib.check(rc, "milnce_forward_shard_peers")
        packs.append(torch.cat([row[r * b:(r + 1) * b], noms[r * b:(r + 1) * b], col]))
        wss.append(ws)
    den, nom, loss = CudaBackend.shard_combine(torch.stack(packs).contiguous(), W, b)
    # the same from W separate buffers (what the ranks' peer-readable statistics packs are)
    den2, nom2, loss2 = torch.empty_like(den), torch.empty_like(nom), torch.empty(1, device="cuda")
    karr = (ctypes.c_void_p * W)(*[x.data_ptr() for x in packs])
    _lib.check(lib.milnce_shard_combine_peers(karr, W, b, _p(den2), _p(nom2), _p(loss2), None, st), "combine_peers")
    assert torch.equal(den, den2) and torch.equal(nom, nom2) and loss.item() == loss2.item()
    # multi-block form (scratch given): same den / nom, loss equal up to the order of the double additions, and
    # the scratch ticket is left at zero so that the call can be repeated
    scratch = torch.zeros(lib.milnce_shard_combine_scratch_bytes(W, b), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        den3, nom3, loss3 = torch.empty_like(den), torch.empty_like(nom), torch.empty(1, device="cuda")
        _lib.check(lib.milnce_shard_combine_peers(karr, W, b, _p(den3), _p(nom3), _p(loss3), _p(scratch), st), "combine_peers")
        assert torch.equal(den, den3) and torch.equal(nom, nom3)
        assert abs(loss.item() - loss3.item()) <= 1e-6 * abs(loss.item())
    assert int(scratch[:4].view(torch.int32)[0]) == 0
    ref_loss, ref_dv, ref_dt, ref_den, ref_nom = milnce_oracle.milnce_closed_form(V.float().cpu().double().numpy(),
                                                                                T.float().cpu().double().numpy(), 1.0)
    assert abs(loss.item() - ref_loss) <= 1e-5 * abs(ref_loss)
    np.testing.assert_allclose(den.cpu().numpy(), ref_den, rtol=2e-5, atol=2e-5)
    np.testing.assert_allclose(nom.cpu().numpy(), ref_nom, rtol=2e-5, atol=2e-5)

    # backward: every "rank" leaves its bf16 partial, then each sums its block of all partials
    gl = torch.ones(1, device="cuda")
    partials = []
    dvs = []
    for r in range(W):
        out = torch.empty(B * K * D, dtype=torch.bfloat16, device="cuda")
        dv, dtp = CudaBackend.backward_shard(V[r * b:(r + 1) * b], T[r * b * K:(r + 1) * b * K], den, nom, gl, r * b, b,
                                             wss[r], parts=3, grad_dtype=torch.float32,
                                             partial_dtype=torch.bfloat16, dims=(B, K, D), partial_out=out)
        assert dtp.data_ptr() == out.data_ptr()
        partials.append(out)
        dvs.append(dv)
    parr = (ctypes.c_void_p * W)(*[x.data_ptr() for x in partials])
    dts = []
    for r in range(W):
        for code, dt_ in ((_lib.F32, torch.float32), (_lib.BF16, torch.bfloat16)):
            dt = torch.empty(b * K, D, dtype=dt_, device="cuda")
            rc = lib.milnce_peer_reduce_dtext(parr, W, r, b, K, D, _p(dt), code, st)
            _lib.check(rc, "milnce_peer_reduce_dtext")
            if code == _lib.F32:
                dts.append(dt)
                # exactly the rank-ordered float32 sum of the bf16 partials
                want = torch.zeros(b * K, D, device="cuda")
                for q in range(W):
                    want += partials[q].view(B * K, D)[r * b * K:(r + 1) * b * K].float()
                assert torch.equal(dt, want)
            else:
                assert torch.equal(dt, dts[-1].bfloat16())
    # push form: every "rank" delivers the rows owned by rank q into q's receive area (block = source rank);
    # summing the W received blocks must give bit-identical gradients (same values, same order)
    block = b * K * D
    recv = [torch.zeros(W * block, dtype=torch.bfloat16, device="cuda") for _ in range(W)]
    for r in range(W):
        dst = (ctypes.c_void_p * W)(*[recv[q].data_ptr() + r * block * 2 for q in range(W)])
        rc = lib.milnce_backward_shard_push(_p(den), _p(nom), _p(gl), dst, W, B, K, D, r * b, b, _p(wss[r]),
                                            wss[r].numel(), st)
        _lib.check(rc, "milnce_backward_shard_push")
    for r in range(W):
        blocks = (ctypes.c_void_p * W)(*[recv[r].data_ptr() + q * block * 2 for q in range(W)])
        dt = torch.empty(b * K, D, device="cuda")
        _lib.check(lib.milnce_peer_reduce_dtext(blocks, W, 0, b, K, D, _p(dt), _lib.F32, st), "reduce (pushed)")
        assert torch.equal(dt, dts[r])
    torch.cuda.synchronize()
    dv = torch.cat(dvs).cpu().double().numpy()
    dt = torch.cat(dts).cpu().double().numpy()
    e_dv = np.linalg.norm(dv - ref_dv) / np.linalg.norm(ref_dv)
    e_dt = np.linalg.norm(dt - ref_dt) / np.linalg.norm(ref_dt)
    assert e_dv < 1e-3, e_dv                      # bf16-input budget
    assert e_dt < 3e-3, e_dt                      # W bf16-rounded partials are summed: one rounding each


def test_peer_reduce_rejects_bad_arguments():
    lib = _lib.load()
    x = torch.zeros(64, dtype=torch.bfloat16, device="cuda")
    arr = (ctypes.c_void_p * 1)(x.data_ptr())
    out = torch.zeros(64, device="cuda")
    assert lib.milnce_peer_reduce_dtext(arr, 17, 0, 1, 1, 64, _p(out), _lib.F32, _lib.stream_ptr()) < 0
    assert lib.milnce_peer_reduce_dtext(arr, 1, 0, 1, 1, 4, _p(out), _lib.F32, _lib.stream_ptr()) < 0   # not % 8
    assert lib.milnce_peer_reduce_dtext(arr, 1, 0, 1, 1, 64, _p(out), 7, _lib.stream_ptr()) < 0
    assert b"milnce_peer_reduce_dtext" in lib.b200head_last_error()
