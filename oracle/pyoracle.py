"""ctypes bindings for the CPU oracle (oracle/libvsk_oracle.so) and, when it
was built, the reference harness (oracle/_ref/libref_harness.so).

TEST INFRASTRUCTURE ONLY -- may be imported from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs, never from vntrseek_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libvsk_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libref_harness.so")

_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build(ref=None):
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref is None:
        ref = os.path.isdir("/root/reference/src")
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


class Oracle:
    """The C restatement (oracle/vsk_oracle.c)."""

    def __init__(self, table):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        L = C.CDLL(ORACLE_SO)
        L.vo_init.argtypes = [_u8p]
        L.vo_dp.argtypes = [_i32p, C.c_int, _i32p, C.c_int, C.c_int, C.c_int, _i32p]
        L.vo_similarity.restype = C.c_double
        L.vo_similarity.argtypes = [C.c_int, C.c_int]
        L.vo_cells.restype = C.c_longlong
        L.vo_cells.argtypes = [C.c_int] * 4
        L.vo_lcs.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int]
        L.vo_lcs_min.argtypes = [C.c_int, C.c_int]
        L.vo_ed.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int]
        L.vo_flank_maxerr.argtypes = [C.c_int]
        L.vo_nbdist.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int]
        L.vo_band.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.vo_cyclic.argtypes = [_i32p, C.c_int, _i32p, C.c_int, _i32p]
        _i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
        _u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
        L.vo_dp_batch.argtypes = [_u16p, _i64p, _u16p, _i64p, _i32p, _i32p, _u8p, C.c_int64, C.c_int, _i32p]
        L.vo_dp_batch.restype = None
        L.vo_lcs_batch.argtypes = [C.c_char_p, _i64p, C.c_char_p, _i64p, _i32p, _i32p, C.c_int64, _i32p]
        L.vo_lcs_batch.restype = None
        L.vo_ed_batch.argtypes = [C.c_char_p, _i64p, _i32p, C.c_char_p, _i64p, _i32p, C.c_int64, C.c_int, _i32p]
        L.vo_ed_batch.restype = None
        L.vo_nbdist_batch.argtypes = [C.c_char_p, _i64p, C.c_char_p, _i64p, C.c_int64, C.c_int, _i32p]
        L.vo_nbdist_batch.restype = None
        self.L = L
        table = np.ascontiguousarray(table, dtype=np.uint8)
        assert table.size == 501501
        L.vo_init(table)

    # batched forms (OpenMP over the pairs): same arithmetic, arrays in / arrays out
    def dp_batch(self, loff, lids, toff, tids, li, ti, mode, homology=88):
        li = np.ascontiguousarray(li, np.int32)
        out = np.zeros((max(len(li), 1), 4), np.int32)
        self.L.vo_dp_batch(np.ascontiguousarray(lids, np.uint16), np.ascontiguousarray(loff, np.int64),
                           np.ascontiguousarray(tids, np.uint16), np.ascontiguousarray(toff, np.int64), li,
                           np.ascontiguousarray(ti, np.int32), np.ascontiguousarray(mode, np.uint8), len(li),
                           homology, out)
        return out[:len(li)]

    def lcs_batch(self, abuf, aoff, bbuf, boff, ia, ib):
        ia = np.ascontiguousarray(ia, np.int32)
        out = np.zeros(max(len(ia), 1), np.int32)
        self.L.vo_lcs_batch(abuf, np.ascontiguousarray(aoff, np.int64), bbuf, np.ascontiguousarray(boff, np.int64),
                            ia, np.ascontiguousarray(ib, np.int32), len(ia), out)
        return out[:len(ia)]

    def ed_batch(self, tbuf, toff, tlen, pbuf, poff, plen, variant):
        n = len(tlen)
        out = np.zeros(max(n, 1), np.int32)
        self.L.vo_ed_batch(tbuf, np.ascontiguousarray(toff, np.int64), np.ascontiguousarray(tlen, np.int32), pbuf,
                           np.ascontiguousarray(poff, np.int64), np.ascontiguousarray(plen, np.int32), n, variant, out)
        return out[:n]

    def nbdist_batch(self, abuf, aoff, bbuf, boff, maxerr):
        n = len(aoff) - 1
        out = np.zeros(max(n, 1), np.int32)
        self.L.vo_nbdist_batch(abuf, np.ascontiguousarray(aoff, np.int64), bbuf, np.ascontiguousarray(boff, np.int64),
                               n, maxerr, out)
        return out[:n]

    def dp(self, left, top, mode=0, homology=88):
        left = np.ascontiguousarray(left, dtype=np.int32)
        top = np.ascontiguousarray(top, dtype=np.int32)
        out = np.zeros(4, np.int32)
        rc = self.L.vo_dp(left, len(left), top, len(top), mode, homology, out)
        assert rc == 0
        return tuple(int(x) for x in out)

    def cyclic(self, left, top):
        """GetBestProfileAP restated: (distance, dsum, sumw, length, start rotation)"""
        left = np.ascontiguousarray(left, dtype=np.int32)
        top = np.ascontiguousarray(top, dtype=np.int32)
        out = np.zeros(5, np.int32)
        assert self.L.vo_cyclic(left, len(left), top, len(top), out) == 0
        return tuple(int(x) for x in out)

    def similarity(self, dsum, sumw):
        return self.L.vo_similarity(dsum, sumw)

    def cells(self, l1, l2, mode=0, homology=88):
        return self.L.vo_cells(l1, l2, mode, homology)

    def band(self, l1, l2, homology=88):
        sv, br = C.c_double(), C.c_int()
        self.L.vo_band(l1, l2, homology, C.byref(sv), C.byref(br))
        return sv.value, br.value

    def lcs(self, s1, s2):
        return self.L.vo_lcs(s1, s2, len(s1), len(s2))

    def lcs_min(self, n, m):
        return self.L.vo_lcs_min(n, m)

    def ed(self, text, pat, variant, N=None, M=None):
        return self.L.vo_ed(text, pat, len(text) if N is None else N,
                            len(pat) if M is None else M, variant)

    def flank_maxerr(self, n):
        return self.L.vo_flank_maxerr(n)

    def nbdist(self, s1, s2, maxerr):
        return self.L.vo_nbdist(s1, s2, len(s1), len(s2), maxerr)


class Reference:
    """The unmodified reference functions (oracle/_ref/libref_harness.so)."""

    def __init__(self, distfile=None):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " (run `make -C oracle ref` where /root/reference exists)")
        L = C.CDLL(REF_SO)
        L.ref_init.argtypes = [C.c_char_p]
        L.ref_dp.restype = C.c_double
        L.ref_dp.argtypes = [_i32p, C.c_int, _i32p, C.c_int, C.c_int, C.c_int, _i32p]
        for f in (L.ref_lcs, L.ref_ed_psearch, L.ref_ed_flankalign):
            f.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int]
        L.ref_nbdist.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int]
        if hasattr(L, "ref_cyclic"):
            L.ref_cyclic.restype = C.c_double
            L.ref_cyclic.argtypes = [_i32p, C.c_int, _i32p, C.c_int, _i32p]
        if hasattr(L, "ref_dp_batch"):
            _i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
            _u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
            L.ref_dp_batch.argtypes = [_u16p, _i64p, _u16p, _i64p, _i32p, _i32p, _u8p, C.c_int64, C.c_int, _i32p]
            L.ref_dp_batch.restype = None
        self.L = L
        distfile = distfile or os.path.join(REF_DIR, "eucledian.dst")
        assert L.ref_init(distfile.encode()) == 0

    def dp(self, left, top, mode=0, homology=88):
        left = np.ascontiguousarray(left, dtype=np.int32)
        top = np.ascontiguousarray(top, dtype=np.int32)
        out = np.zeros(4, np.int32)
        sim = self.L.ref_dp(left, len(left), top, len(top), mode, homology, out)
        return tuple(int(x) for x in out), sim

    def cyclic(self, left, top):
        """the unmodified GetBestProfileAP: ((distance, dsum, sumw, length, start), similarity)"""
        left = np.ascontiguousarray(left, dtype=np.int32)
        top = np.ascontiguousarray(top, dtype=np.int32)
        out = np.zeros(5, np.int32)
        sim = self.L.ref_cyclic(left, len(left), top, len(top), out)
        return tuple(int(x) for x in out), sim

    def dp_batch(self, loff, lids, toff, tids, li, ti, mode, homology=88):
        """the unmodified reference aligners over a pair list, the loop in C (one thread)"""
        li = np.ascontiguousarray(li, np.int32)
        out = np.zeros((max(len(li), 1), 4), np.int32)
        self.L.ref_dp_batch(np.ascontiguousarray(lids, np.uint16), np.ascontiguousarray(loff, np.int64),
                            np.ascontiguousarray(tids, np.uint16), np.ascontiguousarray(toff, np.int64), li,
                            np.ascontiguousarray(ti, np.int32), np.ascontiguousarray(mode, np.uint8), len(li),
                            homology, out)
        return out[:len(li)]

    def lcs(self, s1, s2):
        return self.L.ref_lcs(s1, s2, len(s1), len(s2))

    def ed(self, text, pat, variant, N=None, M=None):
        f = self.L.ref_ed_psearch if variant == 0 else self.L.ref_ed_flankalign
        return f(text, pat, len(text) if N is None else N, len(pat) if M is None else M)

    def nbdist(self, s1, s2, maxerr):
        return self.L.ref_nbdist(s1, s2, len(s1), len(s2), maxerr)


def have_reference():
    return os.path.exists(REF_SO)
