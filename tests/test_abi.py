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


def test_sweep_strip_plan_covers_the_columns():
    """vsk_dd_plan (host side of the sweep's strip arrangement): the strips cover the pair's columns, the last
    strip is the only one that may differ, no arrangement of equal strips is cheaper by the kernel's own cost
    model, and a mixed plan is strictly cheaper than the best equal one"""
    from vntrseek_b200._lib import LIB_PATH
    L = ctypes.CDLL(LIB_PATH)
    L.vsk_dd_plan.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.POINTER(ctypes.c_int32 * 3)]
    cost = lambda c: c * 9 + 41

    def plan(l2, force):
        out = (ctypes.c_int32 * 3)()
        assert L.vsk_dd_plan(l2, force, ctypes.byref(out)) == 0
        return tuple(out)

    mixed = 0
    for l2 in list(range(129, 2200)) + [3000, 4095, 4097, 5999, 6000]:
        cm, nm, cl = plan(l2, 0)
        assert cm in (8, 12, 16) and nm >= 1 and cl in (0, 8, 12, 16) and cl != cm
        cols = nm * 32 * cm + 32 * cl
        assert cols >= l2
        assert cols - l2 < 32 * (cl or cm)           # the last strip is not empty
        if cl:
            assert nm * 32 * cm < l2                 # ... and it is needed
        c0 = nm * cost(cm) + (cost(cl) if cl else 0)
        em, en, ez = plan(l2, 1)
        assert ez == 0 and en == -(-l2 // (32 * em))
        best_equal = min(-(-l2 // (32 * c)) * cost(c) for c in (8, 12, 16))
        assert en * cost(em) == best_equal
        assert c0 <= best_equal and (c0 < best_equal) == (cl != 0)
        mixed += cl != 0
        for f in (8, 12, 16):
            assert plan(l2, f) == (f, -(-l2 // (32 * f)), 0)
    assert mixed > 300
    assert L.vsk_dd_plan(0, 0, ctypes.byref((ctypes.c_int32 * 3)())) != 0
    assert L.vsk_dd_plan(100, 5, ctypes.byref((ctypes.c_int32 * 3)())) != 0
