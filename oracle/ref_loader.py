"""Imports the unmodified reference modules from oracle/_ref (see oracle/make_ref.py).

TEST / BENCH INFRASTRUCTURE ONLY: tests/, __graft_entry__.smoke(), the harness comparisons and
bench.py's `cpu_baseline` / `--impl reference` legs may use this; the product package never does.

The reference hard-codes `.cuda()` on the identity mask of MILNCELoss (loss.py:13) and cannot run
on CPU tensors as written.  `cpu_cuda_shim()` is a context manager under which
`Tensor.cuda()` returns the tensor itself, which is how the reference's own arithmetic is run -
unmodified - on the host cores.
"""
from __future__ import annotations

import contextlib
import importlib
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available() -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in ("loss.py", "soft_dtw_cuda.py", "dtw.py"))


def why_unavailable() -> str:
    if not available():
        return "oracle/_ref is not populated (run python oracle/make_ref.py where /root/reference exists)"
    try:
        import numba  # noqa: F401
        import sklearn  # noqa: F401
    except Exception as e:  # pragma: no cover - depends on the image
        return f"the reference's imports are missing here: {e}"
    return ""


def load(name: str):
    """import <name> from oracle/_ref (loss, soft_dtw_cuda, utils, s3dg)."""
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    mod = sys.modules.get(name)
    if mod is not None and os.path.dirname(os.path.abspath(getattr(mod, "__file__", ""))) != REF_DIR:
        # a module of the same name from somewhere else (the repo's own `utils`, say): do not mix them up
        del sys.modules[name]
    return importlib.import_module(name)


@contextlib.contextmanager
def cpu_cuda_shim():
    import torch
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.Tensor.cuda = orig
