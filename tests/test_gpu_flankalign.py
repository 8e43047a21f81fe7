"""GPU, file level: the drop-in flankalign.exe against the reference flankalign.exe
(oracle/_ref) -- every `<a>_<b>.map` byte for byte -- and the per-pair kernel against
the CPU oracle."""
import filecmp
import os
import shutil

import numpy as np
import pytest

from _cli import OUR_BIN, REF_BIN, run

pytestmark = pytest.mark.gpu


def _same_dirs(a, b):
    fa, fb = sorted(os.listdir(a)), sorted(os.listdir(b))
    assert fa == fb
    for f in fa:
        assert filecmp.cmp(os.path.join(a, f), os.path.join(b, f), shallow=False), f


@pytest.mark.parametrize("nblocks,seed,maxerr,maxflank,paterr", [(40, 5, 0, 50, 15), (25, 6, 3, 20, 30), (60, 7, 0, 1000, 15)])
def test_flankalign_matches_reference(tmp_path, nblocks, seed, maxerr, maxflank, paterr):
    if not os.path.exists(os.path.join(REF_BIN, "flankalign.exe")):
        pytest.skip("oracle/_ref/flankalign.exe not built")
    from vntrseek_b200 import synth
    inp = str(tmp_path / "in.flanks")
    synth.write_flanks(inp, synth.make_flank_blocks(nblocks, nref=8, nread=60, seed=seed))
    a, b = str(tmp_path / "ref_out"), str(tmp_path / "our_out")
    rc1, _, _ = run(os.path.join(REF_BIN, "flankalign.exe"), [a, "xml", inp, maxerr, maxflank, 3, paterr], str(tmp_path))
    rc2, out2, err2 = run(os.path.join(OUR_BIN, "flankalign.exe"), [b, "xml", inp, maxerr, maxflank, 3, paterr], str(tmp_path))
    assert rc1 == 0 and rc2 == 0, (out2[-400:], err2[-400:])
    _same_dirs(a, b)


def test_flankalign_golden_small(tmp_path):
    from conftest import GOLDEN
    g = os.path.join(GOLDEN, "flankalign_small")
    b = str(tmp_path / "our_out")
    rc, out, err = run(os.path.join(OUR_BIN, "flankalign.exe"), [b, "xml", os.path.join(g, "in.flanks"), 0, 50, 1, 15], str(tmp_path))
    assert rc == 0, (out[-400:], err[-400:])
    _same_dirs(os.path.join(g, "expected"), b)


def test_flankalign_stdin_and_errors(tmp_path):
    from conftest import GOLDEN
    exe = os.path.join(OUR_BIN, "flankalign.exe")
    assert run(exe, ["o", "x"], str(tmp_path))[0] == 1                                  # too few arguments
    assert run(exe, ["o", "x", "nofile", 0, 50, 1, 15], str(tmp_path))[0] == 1           # missing input
    assert run(exe, ["o", "x", "nofile", 0, 0, 1, 15], str(tmp_path))[0] == 1            # MAX_FLANK_CONSIDERED < 1
    assert run(exe, ["o", "x", "nofile", 0, 50, 1, 0], str(tmp_path))[0] == 1            # PATLEN_SIZE_ERR < 1


def test_flank_pairs_vs_oracle(ctx, oracle):
    """kernel level, through the C ABI: thresholds, trimming, both orientations"""
    import ctypes as C

    from vntrseek_b200._lib import lib
    rng = np.random.default_rng(3)

    def dna(n):
        return bytes(rng.choice(list(b"ACGTN"), n, p=[.24, .24, .24, .24, .04]).tolist())
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    pool = bytearray()
    recs = {"ref": [], "read": []}
    strs = {"ref": [], "read": []}
    for kind, n in (("ref", 40), ("read", 60)):
        for _ in range(n):
            l, r = dna(int(rng.integers(0, 70))), dna(int(rng.integers(0, 70)))
            if kind == "read" and rng.random() < 0.5 and strs["ref"]:
                src = strs["ref"][int(rng.integers(0, len(strs["ref"])))]
                l, r = src[0][:int(rng.integers(0, 51))], src[1][:int(rng.integers(0, 51))]
                if rng.random() < 0.5:   # opposite strand: left = comp(ref right), right = comp(ref left)
                    l, r = r.translate(comp), l.translate(comp)
            recs[kind].append((len(pool), len(pool) + len(l), len(l), len(r)))
            pool += l + r
            strs[kind].append((l, r))
    dt = np.dtype([("lo", np.int64), ("ro", np.int64), ("ll", np.int32), ("rl", np.int32)])
    refs = np.array(recs["ref"], dtype=dt)
    reads = np.array(recs["read"], dtype=dt)
    pr, pf = np.meshgrid(np.arange(len(reads), dtype=np.int32), np.arange(len(refs), dtype=np.int32), indexing="ij")
    pr, pf = np.ascontiguousarray(pr.ravel()), np.ascontiguousarray(pf.ravel())
    for maxerr in (0, 2):
        out = np.zeros(len(pr), np.uint32)
        ms = C.c_float()
        L = lib()
        L.vsk_flank_pairs.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32,
                                      C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.POINTER(C.c_float)]
        rc = L.vsk_flank_pairs(ctx.h, bytes(pool), len(pool), refs.ctypes.data, len(refs), reads.ctypes.data, len(reads),
                               pr.ctypes.data, pf.ctypes.data, len(pr), maxerr, out.ctypes.data, C.byref(ms))
        assert rc == 0
        for k in range(len(pr)):
            (ql, qr), (fl, fr) = strs["read"][pr[k]], strs["ref"][pf[k]]
            if maxerr:
                m1l = m1r = large = maxerr
            else:
                m1l, m1r = oracle.flank_maxerr(len(ql)), oracle.flank_maxerr(len(qr))
                large = max(m1l, m1r)
            m2l, m2r = (m1r, m1l) if not maxerr else (maxerr, maxerr)
            a, b = min(len(fl), len(ql) + large + 2), min(len(fr), len(qr) + large + 2)
            l1 = oracle.ed(fl, ql, 1, a, min(len(ql), a))
            r1 = oracle.ed(fr, qr, 1, b, min(len(qr), b))
            a, b = min(len(fl), len(qr) + large + 2), min(len(fr), len(ql) + large + 2)
            l2 = oracle.ed(fl, qr.translate(comp), 1, a, min(len(qr), a))
            r2 = oracle.ed(fr, ql.translate(comp), 1, b, min(len(ql), b))
            p1, p2 = (l1 <= m1l and r1 <= m1r), (l2 <= m2l and r2 <= m2r)
            exp = 0
            if p1 or p2:
                if p1 and p2:
                    le, re = (l1, r1) if l1 + r1 <= l2 + r2 else (l2, r2)
                elif p1:
                    le, re = l1, r1
                else:
                    le, re = l2, r2
                exp = le | (re << 8) | (1 << 16)
            assert int(out[k]) == exp, (k, ql, qr, fl, fr)
