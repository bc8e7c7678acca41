"""The C-ABI library builds, loads without a GPU, and exports every symbol that
include/b200head.h declares.  No compute call is made here."""
import ctypes
import os
import re

import pytest

import milnce_b200
from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "b200head.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b([a-z][a-z0-9_]*)\s*\(", text)
    return sorted({n for n in names if n.startswith(("milnce_", "sdtw_", "b200head_"))})


def test_header_declares_expected_entry_points():
    names = _declared()
    for must in ("milnce_forward", "milnce_backward", "milnce_workspace_bytes", "sdtw_forward",
                 "sdtw_backward", "sdtw_aux_bytes", "b200head_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = milnce_b200._lib.load()
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in b200head.h but not exported"
    assert lib.b200head_version() >= 1
    assert isinstance(lib.b200head_last_error(), bytes)


def test_size_queries_are_pure_host_functions():
    lib = milnce_b200._lib.load()
    full = lib.milnce_workspace_bytes(8192, 4, 512, milnce_b200._lib.BF16, 8192)
    part = lib.milnce_workspace_bytes(8192, 4, 512, milnce_b200._lib.BF16, 1024)
    assert 0 < part < full < 4 * 2 ** 30
    assert lib.milnce_workspace_bytes(0, 4, 512, 0, 0) == 0          # rejected shape
    assert lib.milnce_workspace_bytes(64, 4, 512, 7, 64) == 0        # unknown dtype
    assert lib.sdtw_aux_bytes(256, 64, 64, 512) > 256 * 64 * 64 * 4
    assert lib.sdtw_aux_bytes(0, 64, 64, 512) == 0


def test_bad_arguments_are_rejected_before_any_cuda_work():
    lib = milnce_b200._lib.load()
    rc = lib.milnce_forward(None, None, 64, 4, 512, 0, 0, 64, None, None, None, None, 0, None)
    assert rc == -1 and b"null" in lib.b200head_last_error()
    rc = lib.milnce_forward(None, None, 64, 4, 512, 0, 32, 64, None, None, None, None, 0, None)
    assert rc == -1 and b"range" in lib.b200head_last_error()
    rc = lib.sdtw_forward(None, None, 1, 8, 8, 4, ctypes.c_float(1.0), ctypes.c_float(0.0), 0,
                          None, None, None, 0, None)
    assert rc == -1
