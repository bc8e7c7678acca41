"""tcgen05/TMA GEMM core against a plain fp32 torch matmul (GPU).

Validates the TMA tensor maps, the UMMA shared-memory / instruction descriptors,
the smem ring and the TMEM epilogue independently of the loss epilogues.
fp16 products are exact in fp32, so the only difference is accumulation order:
tolerance 2e-6 relative (norm-wise)."""
import ctypes

import pytest
import torch

import milnce_b200
from conftest import rel_err

pytestmark = pytest.mark.gpu


def _gemm(a, b, alpha=1.0):
    lib = milnce_b200._lib.load()
    M, Kd = a.shape
    N = b.shape[0]
    c = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
    rc = lib.b200head_gemm_tn_f16(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                  ctypes.c_void_p(c.data_ptr()), M, N, Kd, a.stride(0), b.stride(0),
                                  c.stride(0), ctypes.c_float(alpha), milnce_b200._lib.stream_ptr())
    milnce_b200._lib.check(rc, "gemm_tn_f16")
    torch.cuda.synchronize()
    return c


@pytest.mark.parametrize("M,N,Kd", [
    (128, 128, 64),      # one tile, one k-block
    (128, 128, 512),     # k loop, ring wraps
    (256, 384, 1024),    # several tiles per CTA? no: 6 tiles on 6 CTAs
    (8192, 512, 256),    # more tiles than SMs: persistent loop + TMEM double buffer
    (100, 200, 72),      # ragged M, N and K (TMA zero fill)
    (1, 8, 8),           # degenerate
    (300, 130, 1544),    # K not a multiple of 64, N just over one tile
])
def test_gemm_tn_matches_torch(M, N, Kd):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + Kd)
    lda = (Kd + 7) // 8 * 8
    a_full = torch.randn(M, lda, generator=g, device="cuda").half()
    b_full = torch.randn(N, lda, generator=g, device="cuda").half()
    a, b = a_full[:, :Kd], b_full[:, :Kd]
    c = _gemm(a, b, alpha=0.5)
    ref = 0.5 * (a.float() @ b.float().t())
    assert not torch.isnan(c).any()
    err = rel_err(c.cpu().numpy(), ref.cpu().numpy())
    print(f"gemm {M}x{N}x{Kd}: rel err {err:.3e}")
    assert err < 2e-6


def test_gemm_identity_pattern_localises_descriptor_bugs():
    """A = one-hot rows, B = index-coded columns: every output names its own (m, n)."""
    M = N = 128
    Kd = 64
    a = torch.zeros(M, Kd, device="cuda", dtype=torch.float16)
    a[torch.arange(M), torch.arange(M) % Kd] = 1.0
    b = (torch.arange(N, device="cuda").float()[:, None] * 64 +
         torch.arange(Kd, device="cuda").float()[None, :]).half()
    c = _gemm(a, b)
    ref = a.float() @ b.float().t()
    bad = (c != ref).nonzero()
    assert bad.numel() == 0, f"first mismatches (m, n): {bad[:8].tolist()}"


def _gemm_at(at, b, alpha=1.0):
    """A passed transposed: at is (Kd, M).  Exercises the MN-major A descriptor."""
    lib = milnce_b200._lib.load()
    Kd, M = at.shape
    N = b.shape[0]
    c = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
    rc = lib.b200head_gemm_at_f16(ctypes.c_void_p(at.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                  ctypes.c_void_p(c.data_ptr()), M, N, Kd, at.stride(0), b.stride(0),
                                  c.stride(0), ctypes.c_float(alpha), milnce_b200._lib.stream_ptr())
    milnce_b200._lib.check(rc, "gemm_at_f16")
    torch.cuda.synchronize()
    return c


@pytest.mark.parametrize("M,N,Kd", [
    (128, 128, 64),
    (128, 128, 256),
    (256, 256, 1024),
    (4096, 512, 2048),   # the dtext GEMM shape class (M = B*K, N = D, K = B)
    (200, 72, 136),      # ragged everywhere
])
def test_gemm_mn_major_a_matches_torch(M, N, Kd):
    g = torch.Generator(device="cuda").manual_seed(M + N * 5 + Kd * 11)
    ldat = (M + 7) // 8 * 8
    ldb = (Kd + 7) // 8 * 8
    at = torch.randn(Kd, ldat, generator=g, device="cuda").half()[:, :M]
    b = torch.randn(N, ldb, generator=g, device="cuda").half()[:, :Kd]
    c = _gemm_at(at, b, alpha=2.0)
    ref = 2.0 * (at.float().t() @ b.float().t())
    assert not torch.isnan(c).any()
    err = rel_err(c.cpu().numpy(), ref.cpu().numpy())
    print(f"gemm_at {M}x{N}x{Kd}: rel err {err:.3e}")
    assert err < 2e-6


@pytest.mark.parametrize("transposed", [0, 1, 2, 3])
@pytest.mark.parametrize("M,N,Kd", [
    (256, 256, 64),      # one pair tile, one k-block
    (256, 256, 512),
    (1024, 512, 2048),   # several tiles per pair: TMEM double buffering across tiles
    (8192, 512, 256),    # more tiles than CTA pairs
    (300, 200, 136),     # ragged: the follower CTA's rows / columns partly or fully out of range
    (100, 72, 72),
])
def test_gemm_cta_pair_kernel_matches_torch(M, N, Kd, transposed):
    """tcgen05.mma.cta_group::2 path: 2-CTA clusters, peer-signalled TMA, multicast commits.
    transposed bit 0: A passed as (Kd, M) (MN-major A); bit 1: B passed as (Kd, N) (MN-major B)."""
    lib = milnce_b200._lib.load()
    g = torch.Generator(device="cuda").manual_seed(M * 3 + N * 7 + Kd + transposed)
    ldk = (Kd + 7) // 8 * 8
    if transposed & 2:
        ldn = (N + 7) // 8 * 8
        b = torch.randn(Kd, ldn, generator=g, device="cuda").half()[:, :N]
        bt = b.float()                    # (Kd, N)
    else:
        b = torch.randn(N, ldk, generator=g, device="cuda").half()[:, :Kd]
        bt = b.float().t()                # (Kd, N)
    if transposed & 1:
        ldm = (M + 7) // 8 * 8
        a = torch.randn(Kd, ldm, generator=g, device="cuda").half()[:, :M]
        ref = a.float().t() @ bt
    else:
        a = torch.randn(M, ldk, generator=g, device="cuda").half()[:, :Kd]
        ref = a.float() @ bt
    c = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
    rc = lib.b200head_gemm2_f16(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                ctypes.c_void_p(c.data_ptr()), M, N, Kd, a.stride(0), b.stride(0),
                                c.stride(0), ctypes.c_float(1.0), transposed,
                                milnce_b200._lib.stream_ptr())
    milnce_b200._lib.check(rc, "gemm2_f16")
    torch.cuda.synchronize()
    assert not torch.isnan(c).any()
    err = rel_err(c.cpu().numpy(), ref.cpu().numpy())
    print(f"gemm2 {M}x{N}x{Kd} t={transposed}: rel err {err:.3e}")
    assert err < 3e-6


@pytest.mark.parametrize("transposed", [0, 2, 3])
def test_gemm_cta_pair_kernel_bf16_operands(transposed):
    """Bit 2 of the flag: both operands are bfloat16 (the instruction descriptor's A and B format
    fields set to BF16).  Mixed fp16 x bf16 operands are not offered: the hardware rejects them."""
    lib = milnce_b200._lib.load()
    M, N, Kd = 512, 256, 320
    g = torch.Generator(device="cuda").manual_seed(40 + transposed)
    b = torch.randn((Kd, N) if transposed & 2 else (N, Kd), generator=g, device="cuda").bfloat16()
    bt = b.float() if transposed & 2 else b.float().t()
    a = torch.randn((Kd, M) if transposed & 1 else (M, Kd), generator=g, device="cuda").bfloat16()
    ref = (a.float().t() if transposed & 1 else a.float()) @ bt
    c = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
    rc = lib.b200head_gemm2_f16(ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                ctypes.c_void_p(c.data_ptr()), M, N, Kd, a.stride(0), b.stride(0),
                                c.stride(0), ctypes.c_float(1.0), transposed | 4,
                                milnce_b200._lib.stream_ptr())
    milnce_b200._lib.check(rc, "gemm2_f16")
    torch.cuda.synchronize()
    err = rel_err(c.cpu().numpy(), ref.cpu().numpy())
    print(f"gemm2 bf16 operands t={transposed}: rel err {err:.3e}")
    assert err < 3e-6
