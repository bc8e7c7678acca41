"""Out-of-bounds canaries (GPU): every buffer handed to the C ABI sits between guard regions
that must come back untouched.  Complements the parity tests: a kernel that writes past an
edge tile, a ragged row or its workspace slice usually still produces the right numbers."""
import ctypes

import pytest
import torch

import milnce_b200
from milnce_b200 import _lib

pytestmark = pytest.mark.gpu

GUARD = 4096          # bytes on either side
PATTERN = 0xA5


class Guarded:
    """A device buffer of `nbytes` (256-byte aligned start) with guard bytes before and after."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        pad = (256 - self.nbytes % 256) % 256
        self.raw = torch.full((GUARD + self.nbytes + pad + GUARD,), PATTERN, dtype=torch.uint8, device="cuda")
        self.body = self.raw[GUARD:GUARD + self.nbytes]
        assert self.body.data_ptr() % 256 == 0

    def view(self, dtype, shape):
        return self.body.view(dtype).view(shape)

    def ptr(self):
        return ctypes.c_void_p(self.body.data_ptr())

    def intact(self):
        head = self.raw[:GUARD]
        tail = self.raw[GUARD + self.nbytes:]
        return bool((head == PATTERN).all()) and bool((tail == PATTERN).all())


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
# (64, 4, 100) and (96, 2, 36): D not a multiple of 8 - bf16 inputs then leave the vectorised staging
# kernels for the generic ones (which once wrote transposed copies the one-plane layout has no room for)
@pytest.mark.parametrize("B,K,D", [(300, 4, 136), (129, 3, 72), (64, 1, 8), (520, 2, 264), (64, 4, 100),
                                   (96, 2, 36)])
def test_milnce_stays_inside_its_buffers(B, K, D, dtype):
    lib = _lib.load()
    code = _lib.BF16 if dtype == torch.bfloat16 else _lib.F32
    es = 2 if dtype == torch.bfloat16 else 4
    g = torch.Generator().manual_seed(B + K + D)
    bufs = {
        "video": Guarded(B * D * es), "text": Guarded(B * K * D * es),
        "den": Guarded(B * 4), "nom": Guarded(B * 4), "loss": Guarded(4),
        "dv": Guarded(B * D * 4), "dt": Guarded(B * K * D * 4), "gl": Guarded(4),
        "ws": Guarded(lib.milnce_workspace_bytes(B, K, D, code, B)),
    }
    bufs["video"].view(dtype, (B, D)).copy_((torch.randn(B, D, generator=g) * 0.3).to(dtype))
    bufs["text"].view(dtype, (B * K, D)).copy_((torch.randn(B * K, D, generator=g) * 0.3).to(dtype))
    bufs["gl"].view(torch.float32, (1,)).fill_(1.0)
    st = _lib.stream_ptr()
    rc = lib.milnce_forward(bufs["video"].ptr(), bufs["text"].ptr(), B, K, D, code, 0, B, bufs["den"].ptr(),
                            bufs["nom"].ptr(), bufs["loss"].ptr(), bufs["ws"].ptr(), bufs["ws"].nbytes, st)
    _lib.check(rc, "milnce_forward")
    rc = lib.milnce_backward(bufs["video"].ptr(), bufs["text"].ptr(), bufs["den"].ptr(), bufs["nom"].ptr(),
                             bufs["gl"].ptr(), bufs["dv"].ptr(), bufs["dt"].ptr(), _lib.F32, B, K, D, code, 0, B, 1,
                             bufs["ws"].ptr(), bufs["ws"].nbytes, st)
    _lib.check(rc, "milnce_backward")
    torch.cuda.synchronize()
    assert torch.isfinite(bufs["loss"].view(torch.float32, (1,))).all()
    for name, b in bufs.items():
        assert b.intact(), f"guard bytes around `{name}` were overwritten"


@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("Bp,N,M,d", [(5, 64, 64, 128), (3, 37, 50, 36), (2, 8, 6, 4), (2, 100, 70, 20)])
def test_sdtw_stays_inside_its_buffers(Bp, N, M, d, kind):
    lib = _lib.load()
    g = torch.Generator().manual_seed(N * 7 + M + d + kind)
    aux_bytes = lib.sdtw_aux_bytes(Bp, N, M, d)
    bufs = {
        "x": Guarded(Bp * N * d * 4), "y": Guarded(Bp * M * d * 4), "R": Guarded(Bp * (N + 2) * (M + 2) * 4),
        "out": Guarded(Bp * 4), "aux": Guarded(aux_bytes), "go": Guarded(Bp * 4), "E": Guarded(Bp * N * M * 4),
        "dx": Guarded(Bp * N * d * 4), "dy": Guarded(Bp * M * d * 4),
    }
    bufs["x"].view(torch.float32, (Bp, N, d)).copy_(torch.randn(Bp, N, d, generator=g) * 0.5)
    bufs["y"].view(torch.float32, (Bp, M, d)).copy_(torch.randn(Bp, M, d, generator=g) * 0.5)
    bufs["go"].view(torch.float32, (Bp,)).fill_(1.0)
    st = _lib.stream_ptr()
    rc = lib.sdtw_forward(bufs["x"].ptr(), bufs["y"].ptr(), Bp, N, M, d, ctypes.c_float(0.5), ctypes.c_float(0.0),
                          kind, bufs["R"].ptr(), bufs["out"].ptr(), bufs["aux"].ptr(), aux_bytes, st)
    _lib.check(rc, "sdtw_forward")
    rc = lib.sdtw_backward(bufs["x"].ptr(), bufs["y"].ptr(), bufs["go"].ptr(), Bp, N, M, d, ctypes.c_float(0.5),
                           ctypes.c_float(0.0), kind, bufs["aux"].ptr(), aux_bytes, bufs["E"].ptr(),
                           bufs["dx"].ptr(), bufs["dy"].ptr(), st)
    _lib.check(rc, "sdtw_backward")
    torch.cuda.synchronize()
    assert torch.isfinite(bufs["out"].view(torch.float32, (Bp,))).all()
    for name, b in bufs.items():
        assert b.intact(), f"guard bytes around `{name}` were overwritten"
