"""CPU: libvsk.so loads and exports every symbol include/vsk.h declares; no compute."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "vsk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vsk_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    assert "vsk_init" in names and "vsk_pair_dp" in names and len(names) >= 10


def test_library_exports_every_declared_symbol():
    from vntrseek_b200._lib import LIB_PATH
    assert os.path.exists(LIB_PATH), "build with __graft_entry__.build()"
    L = ctypes.CDLL(LIB_PATH)
    missing = [n for n in _declared() if not hasattr(L, n)]
    assert not missing, missing


def test_fails_loudly_without_gpu(table):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import vntrseek_b200 as v
    with pytest.raises(v.VskError):
        v.Context(table)


def test_rejects_bad_table():
    import vntrseek_b200 as v
    with pytest.raises(v.VskError):
        v.Context(np.zeros(10, np.uint8))
