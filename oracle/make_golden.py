"""Mint golden vectors from the EXECUTED reference.  Authoring-container only.

Run as ``python oracle/make_golden.py`` where ``/root/reference`` exists.  It
imports the reference's ``loss.py`` and ``soft_dtw_cuda.py`` unmodified, runs
them on seeded synthetic inputs and writes small ``.npz`` fixtures under
``tests/golden/``.  The fixtures travel to the GPU box; the reference does not.

The reference ships no golden vectors or known-answer tests of its own
(SURVEY.md section 4), so these fixtures are what pins the oracle.

Two deviations needed to execute the reference on a CPU-only machine, neither
of which touches its arithmetic:
* ``torch.Tensor.cuda`` is replaced by the identity while ``MILNCELoss`` runs,
  because ``loss.py:13`` moves the identity mask to the GPU unconditionally;
* ``SoftDTW`` is constructed with ``use_cuda=False`` (the numba-CPU path, which
  is the float64-internal one - SURVEY.md F4/F5).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("MILNCE_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _import_reference():
    sys.path.insert(0, REF)
    import loss as ref_loss  # noqa: E402
    import soft_dtw_cuda as ref_sdtw  # noqa: E402
    return ref_loss, ref_sdtw


def golden_milnce(ref_loss):
    cases = [
        # name, B, K, D, scale, seed
        ("tiny", 8, 4, 32, 0.5, 11),
        ("k1", 6, 1, 16, 0.7, 12),
        ("k3_odd", 10, 3, 40, 0.3, 13),
        ("cfg1_b64", 64, 4, 512, 0.05, 0),   # BASELINE.json configs[0]
        ("wide_range", 16, 4, 64, 1.5, 14),  # logits spread over +-40
    ]
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for name, b, k, d, scale, seed in cases:
            g = torch.Generator().manual_seed(seed)
            v32 = (torch.randn(b, d, generator=g) * scale)
            t32 = (torch.randn(b * k, d, generator=g) * scale)
            v = v32.double().requires_grad_(True)
            t = t32.double().requires_grad_(True)
            out = ref_loss.MILNCELoss()(v, t)
            out.backward()
            # the reference evaluated in float32 as well, for a noise-floor record
            v_f = v32.clone().requires_grad_(True)
            t_f = t32.clone().requires_grad_(True)
            out32 = ref_loss.MILNCELoss()(v_f, t_f)
            out32.backward()
            big = b * k * d > 50000  # keep the fixture small: fp32 copies of the f64 grads
            gdt = np.float32 if big else np.float64
            extra = {} if big else dict(dvideo_ref_fp32=v_f.grad.numpy(),
                                        dtext_ref_fp32=t_f.grad.numpy())
            np.savez_compressed(
                os.path.join(OUT, f"milnce_{name}.npz"),
                video=v32.numpy(), text=t32.numpy(), K=np.int64(k),
                loss=np.float64(out.item()),
                dvideo=v.grad.numpy().astype(gdt), dtext=t.grad.numpy().astype(gdt),
                loss_ref_fp32=np.float32(out32.item()), **extra,
            )
            print(f"milnce_{name}: loss={out.item():.12f}")
    finally:
        torch.Tensor.cuda = orig_cuda


def golden_sdtw(ref_sdtw):
    shapes = [
        # name, B, N, M, d, bandwidth, seed, kind
        ("sq8", 4, 8, 8, 16, None, 21, "randn"),
        ("n8m6", 3, 8, 6, 16, None, 22, "randn"),
        ("n6m8", 3, 6, 8, 16, None, 23, "randn"),
        ("n7m5_bw2", 2, 7, 5, 16, 2, 24, "randn"),
        ("sq8_bw2", 2, 8, 8, 16, 2, 25, "randn"),
        ("rand12", 2, 12, 12, 24, None, 26, "rand"),  # torch.rand, like profile() :427-428
    ]
    for name, b, n, m, d, bw, seed, kind in shapes:
        g = torch.Generator().manual_seed(seed)
        gen = torch.randn if kind == "randn" else torch.rand
        x32 = gen(b, n, d, generator=g)
        y32 = gen(b, m, d, generator=g)
        if name == "sq8":
            x32[1, 3, :] = 0.0  # a zero frame: cosine clamp path (eps=1e-8)
        go32 = torch.rand(b, generator=g) + 0.5
        rec = dict(x=x32.numpy(), y=y32.numpy(), grad_out=go32.numpy(),
                   bandwidth=np.float64(0 if bw is None else bw))
        for dist in ("cosine", "negative_dot", "euclidean"):
            for gamma in (0.1, 1.0):
                for normalize in ((False, True) if n == m else (False,)):
                    x = x32.double().requires_grad_(True)
                    y = y32.double().requires_grad_(True)
                    mod = ref_sdtw.SoftDTW(False, gamma=gamma, normalize=normalize,
                                           bandwidth=bw, dist_func=dist)
                    out = mod(x, y)
                    (out * go32.double()).sum().backward()
                    key = f"{dist}_g{gamma}_n{int(normalize)}"
                    rec[key + "_out"] = out.detach().numpy()
                    rec[key + "_dx"] = x.grad.numpy()
                    rec[key + "_dy"] = y.grad.numpy()
                    if not normalize:
                        # direct calls with float64 D: the untruncated recurrences
                        D = mod.dist_func(x32.double(), y32.double()).numpy()
                        g32 = float(np.float32(gamma))
                        bwf = float(0 if bw is None else bw)
                        R = ref_sdtw.compute_softdtw(D, g32, bwf)
                        rec[key + "_R"] = R.copy()
                        rec[key + "_E"] = ref_sdtw.compute_softdtw_backward(D, R.copy(), g32, bwf)
                        rec[key + "_D"] = D
        np.savez_compressed(os.path.join(OUT, f"sdtw_{name}.npz"), **rec)
        print(f"sdtw_{name}: {len(rec)} arrays")


def golden_composite(ref_loss, ref_sdtw):
    """The composite soft-DTW losses of reference loss.py:20-134 (CDTW, SDTW_CIDM, SDTW_negative, SDTW_3), executed with
    the one deviation the docstring above names: every ``SoftDTW(use_cuda=True, ...)`` they construct is
    built with ``use_cuda=False`` (numba-CPU path); their own code is not touched."""
    class Args:
        rank = 1

    real = ref_sdtw.SoftDTW

    def cpu_sdtw(use_cuda, *a, **k):
        return real(False, *a, **k)

    ref_loss.SoftDTW = cpu_sdtw
    try:
        rec = {}
        g = torch.Generator().manual_seed(31)
        # CDTW: batch of 8 (the reference hard-codes repeat(8, ...), :29), frames 6 / words 5, d = 16
        v32 = torch.randn(8, 6, 16, generator=g)
        t32 = torch.randn(8, 5, 16, generator=g)
        v = v32.double().requires_grad_(True)
        t = t32.double().requires_grad_(True)
        out = ref_loss.CDTW(Args())(v, t)
        out.sum().backward()
        rec.update(cdtw_video=v32.numpy(), cdtw_text=t32.numpy(), cdtw_out=out.detach().numpy(),
                   cdtw_dvideo=v.grad.numpy(), cdtw_dtext=t.grad.numpy())
        # SDTW_3: b = 4 sequences of 6 frames, d = 16 (all 16 pairs per term)
        v32 = torch.randn(4, 6, 16, generator=g) * 0.5
        t32 = torch.randn(4, 6, 16, generator=g) * 0.5
        v = v32.double().requires_grad_(True)
        t = t32.double().requires_grad_(True)
        l1, l2, l3 = ref_loss.SDTW_3(Args())(v, t)
        (l1 + 2.0 * l2 + 3.0 * l3).backward()
        rec.update(s3_video=v32.numpy(), s3_text=t32.numpy(), s3_losses=np.array([l1.item(), l2.item(), l3.item()]),
                   s3_dvideo=v.grad.numpy(), s3_dtext=t.grad.numpy())
        # SDTW_CIDM: b = 3, n = m = 6, start times spread so that both branches of `y` occur
        v32 = torch.randn(3, 6, 16, generator=g)
        t32 = torch.randn(3, 6, 16, generator=g)
        start = torch.cumsum(torch.rand(3, 6, generator=g) * 8.0, dim=1).double()
        end = start + 1.0
        v = v32.double().requires_grad_(True)
        t = t32.double().requires_grad_(True)
        out = ref_loss.SDTW_CIDM(Args())(v, t, start, end)
        out.backward()
        rec.update(cidm_video=v32.numpy(), cidm_text=t32.numpy(), cidm_start=start.numpy(), cidm_end=end.numpy(),
                   cidm_out=np.float64(out.item()), cidm_dvideo=v.grad.numpy(), cidm_dtext=t.grad.numpy())
        # SDTW_negative: the reference hard-codes 160 sequences of 8 frames, d = 512 (:80-88), so the fixture
        # has that size; inputs are multiples of 1/32 stored as int8, gradients kept for two sequences only
        vi = torch.randint(-8, 9, (160, 8, 512), generator=g, dtype=torch.int8)
        ti = torch.randint(-8, 9, (160, 8, 512), generator=g, dtype=torch.int8)
        v = (vi.double() / 32).requires_grad_(True)
        t = (ti.double() / 32).requires_grad_(True)
        out = ref_loss.SDTW_negative(Args())(v, t)
        out.backward()
        rec.update(neg_video_i8=vi.numpy(), neg_text_i8=ti.numpy(), neg_out=np.float64(out.item()),
                   neg_dvideo_first2=v.grad[:2].numpy(), neg_dtext_first2=t.grad[:2].numpy(),
                   neg_dvideo_norm=np.float64(v.grad.norm().item()), neg_dtext_norm=np.float64(t.grad.norm().item()))
        np.savez_compressed(os.path.join(OUT, "composite_losses.npz"), **rec)
        print(f"composite_losses: {len(rec)} arrays; CDTW {rec['cdtw_out']}, SDTW_3 {rec['s3_losses']}, "
              f"SDTW_CIDM {rec['cidm_out']}, SDTW_negative {rec['neg_out']}")
    finally:
        ref_loss.SoftDTW = real


def main():
    os.makedirs(OUT, exist_ok=True)
    ref_loss, ref_sdtw = _import_reference()
    if "--composite-only" not in sys.argv:
        golden_milnce(ref_loss)
        golden_sdtw(ref_sdtw)
    golden_composite(ref_loss, ref_sdtw)


if __name__ == "__main__":
    main()
