"""Host-side mirror of the reference module API: constructor attributes, argument
validation and error behaviour.  No GPU needed (every path here raises before any
device work)."""
import pytest
import torch

import milnce_b200
from milnce_b200 import MILNCELoss, SoftDTW


def test_softdtw_attributes_follow_reference():
    m = SoftDTW(True, gamma=0.1, normalize=True, bandwidth=None, dist_func="cosine")
    assert m.gamma == 0.1 and m.normalize is True and m.use_cuda is True
    assert m.bandwidth == 0 and m.dist_func == "cosine"
    assert SoftDTW(True, bandwidth=3, dist_func="euclidean").bandwidth == 3.0


def test_softdtw_dist_func_quirks_are_mirrored():
    # reference soft_dtw_cuda.py:297-304: None leaves the attribute unset -> AttributeError at forward
    m = SoftDTW(True, gamma=1.0)
    assert not hasattr(m, "dist_func")
    with pytest.raises(AttributeError):
        SoftDTW(True, dist_func="negative_cosine")


def test_softdtw_validation_errors():
    m = SoftDTW(True, dist_func="cosine")
    x = torch.zeros(2, 4, 8)
    with pytest.raises(ValueError):
        m(x, torch.zeros(3, 4, 8))          # batch mismatch (reference assert :313)
    with pytest.raises(ValueError):
        m(x, torch.zeros(2, 4, 7))          # feature mismatch (reference assert :314)
    with pytest.raises(RuntimeError):
        SoftDTW(False, dist_func="cosine")(x, x)   # no CPU path
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 1025, 2), torch.zeros(1, 4, 2))  # > 1024: reference falls back to CPU
    with pytest.raises(RuntimeError):
        m(x, x)                              # CPU tensors


def test_milnce_validation_errors():
    loss = MILNCELoss()
    with pytest.raises(ValueError):
        loss(torch.zeros(4, 8), torch.zeros(10, 8))   # rows not a multiple (reference view() fails)
    with pytest.raises(ValueError):
        loss(torch.zeros(4, 8), torch.zeros(8, 9))
    with pytest.raises(TypeError):
        loss(torch.zeros(4, 8), torch.zeros(8, 8, dtype=torch.bfloat16))
    with pytest.raises(RuntimeError):
        loss(torch.zeros(4, 8), torch.zeros(8, 8))    # CPU tensors: no fallback


def test_no_oracle_import_in_product_path():
    """The shipped package must never import the CPU oracle."""
    import os
    pkg = os.path.dirname(milnce_b200.__file__)
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, fn)).read(), fn
