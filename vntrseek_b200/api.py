"""Thin Python mirror of the libvsk C ABI (include/vsk.h).  Every call goes
through the shared library; nothing here computes alignment results."""
import ctypes as C

import numpy as np

from ._lib import lib


class VskError(RuntimeError):
    pass


def pack_profiles(profiles):
    """list of int sequences -> (off int64[n+1], ids uint16[total])"""
    lens = np.fromiter((len(p) for p in profiles), dtype=np.int64, count=len(profiles))
    off = np.zeros(len(profiles) + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    ids = np.zeros(max(int(off[-1]), 1), np.uint16)
    for i, p in enumerate(profiles):
        ids[off[i]:off[i + 1]] = p
    return off, ids


def pack_strings(strings):
    """list of bytes -> (buf bytes, off int64[n+1])"""
    lens = np.fromiter((len(s) for s in strings), dtype=np.int64, count=len(strings))
    off = np.zeros(len(strings) + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    return b"".join(strings), off


class ProfileSet:
    def __init__(self, ctx, off, ids):
        self.ctx = ctx
        self.off = np.ascontiguousarray(off, np.int64)
        self.ids = np.ascontiguousarray(ids, np.uint16)
        self.n = len(self.off) - 1
        self.h = C.c_void_p()
        ctx._check(lib().vsk_profset_upload(ctx.h, self.n, self.off, self.ids, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().vsk_profset_free(self.ctx.h, self.h)
            self.h = C.c_void_p()


class DpBatch:
    def __init__(self, ctx, left, top, li, ti, mode, homology=88):
        self.ctx = ctx
        self.n = len(li)
        self.li = np.ascontiguousarray(li, np.int32)
        self.ti = np.ascontiguousarray(ti, np.int32)
        self.mode = np.ascontiguousarray(mode, np.uint8)
        self.h = C.c_void_p()
        ctx._check(lib().vsk_dp_batch_prepare(ctx.h, left.h, top.h, self.n, self.li, self.ti,
                                              self.mode, homology, C.byref(self.h)))

    @property
    def cells(self):
        return lib().vsk_dp_batch_cells(self.h)

    def run(self):
        self.ctx._check(lib().vsk_dp_batch_run(self.ctx.h, self.h))

    def fetch(self):
        out = np.zeros((max(self.n, 1), 4), np.int32)
        self.ctx._check(lib().vsk_dp_batch_fetch(self.ctx.h, self.h, out))
        return out[:self.n]

    def close(self):
        if self.h:
            lib().vsk_dp_batch_free(self.ctx.h, self.h)
            self.h = C.c_void_p()


class Context:
    """vsk_ctx: one per process/GPU. `table` = the 501,501 distance bytes."""

    def __init__(self, table, device=0):
        table = np.ascontiguousarray(table, np.uint8)
        if table.size != 501501:
            raise VskError("distance table must have 501501 bytes")
        self.h = C.c_void_p()
        rc = lib().vsk_init(C.byref(self.h), device, table)
        if rc != 0:
            raise VskError(f"vsk_init failed ({rc}): no usable CUDA device or bad table")

    def _check(self, rc):
        if rc != 0:
            raise VskError(f"libvsk error {rc}: {lib().vsk_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            lib().vsk_destroy(self.h)
            self.h = C.c_void_p()

    def sync(self):
        self._check(lib().vsk_sync(self.h))

    @property
    def launches(self):
        return lib().vsk_launch_count(self.h)

    def profiles(self, profiles=None, off=None, ids=None):
        if profiles is not None:
            off, ids = pack_profiles(profiles)
        return ProfileSet(self, off, ids)

    def pair_dp(self, left, top, li, ti, mode, homology=88):
        li = np.ascontiguousarray(li, np.int32)
        ti = np.ascontiguousarray(ti, np.int32)
        mode = np.ascontiguousarray(mode, np.uint8)
        out = np.zeros((max(len(li), 1), 4), np.int32)
        self._check(lib().vsk_pair_dp(self.h, left.h, top.h, len(li), li, ti, mode, homology, out))
        return out[:len(li)]

    def dp_batch(self, left, top, li, ti, mode, homology=88):
        return DpBatch(self, left, top, li, ti, mode, homology)

    def lcs_pairs(self, a, b, ia=None, ib=None):
        """a, b: lists of bytes; LCS of a[ia[k]] vs b[ib[k]] as psearch computes it."""
        abuf, aoff = pack_strings(a)
        bbuf, boff = pack_strings(b)
        if ia is None:
            ia = np.arange(len(a), dtype=np.int32)
            ib = np.arange(len(b), dtype=np.int32)
        ia = np.ascontiguousarray(ia, np.int32)
        ib = np.ascontiguousarray(ib, np.int32)
        out = np.zeros(max(len(ia), 1), np.int32)
        self._check(lib().vsk_lcs_pairs(self.h, abuf, aoff, bbuf, boff, len(ia), ia, ib, out))
        return out[:len(ia)]

    def ed_pairs(self, texts, pats, variant, tlen=None, plen=None):
        tbuf, toff = pack_strings(texts)
        pbuf, poff = pack_strings(pats)
        n = len(texts)
        if tlen is None:
            tlen = np.diff(toff)
        if plen is None:
            plen = np.diff(poff)
        tlen = np.ascontiguousarray(tlen, np.int32)
        plen = np.ascontiguousarray(plen, np.int32)
        out = np.zeros(max(n, 1), np.int32)
        self._check(lib().vsk_ed_pairs(self.h, tbuf, np.ascontiguousarray(toff[:-1]), tlen, pbuf,
                                       np.ascontiguousarray(poff[:-1]), plen, n, variant, out))
        return out[:n]
