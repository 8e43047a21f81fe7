"""Thin Python mirror of the libvsk C ABI (include/vsk.h).  Every call goes
through the shared library; nothing here computes alignment results."""
import ctypes as C
import os

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
        self.device = device
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

    def timer_start(self):
        self._check(lib().vsk_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self._check(lib().vsk_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def profile(self, on):
        self._check(lib().vsk_profile_enable(self.h, int(on)))

    def profile_read(self):
        ms, n = C.c_double(), C.c_int64()
        self._check(lib().vsk_profile_read(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def profile_read_sweep(self):
        """(summed ms, launches, busy ms) of the forward-sweep kernel alone since the last call; busy =
        time during which at least one of its launches was running (they overlap across two streams)"""
        ms, n, busy = C.c_double(), C.c_int64(), C.c_double()
        self._check(lib().vsk_profile_read_sweep(self.h, C.byref(ms), C.byref(n), C.byref(busy)))
        return ms.value, n.value, busy.value

    def profile_stages(self):
        buf = C.create_string_buffer(4096)
        self._check(lib().vsk_profile_stages(self.h, buf, 4096))
        return {k: float(v) for k, v in (kv.split("=") for kv in buf.value.decode().split(";") if kv)}

    def int_peak_gops(self):
        g = C.c_double()
        self._check(lib().vsk_int_peak(self.h, C.byref(g)))
        return g.value

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

    def pair_dp_cyclic(self, left, top, li, ti):
        """vsk_pair_dp_cyclic (GetBestProfileAP): rows of (distance, dsum, sumw, length, start rotation)"""
        li = np.ascontiguousarray(li, np.int32)
        ti = np.ascontiguousarray(ti, np.int32)
        out = np.zeros((max(len(li), 1), 8), np.int32)
        self._check(lib().vsk_pair_dp_cyclic(self.h, left.h, top.h, len(li), li, ti, out))
        return out[:len(li), :5]

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


# ---------------------------------------------------------------------------
# PROCLU: TR sets, reference index, per-read scan, clustering
# ---------------------------------------------------------------------------
from ._lib import LINK_DTYPE, ProcluParams, ProcluStats  # noqa: E402


class TRSetHandle:
    """vsk_trset: one .leb36 file in host memory (vsk_trset_load_leb36 / _from_arrays)."""

    def __init__(self, h):
        self.h = h

    @classmethod
    def from_file(cls, path, is_read, maxflank=50, option_r=0, reflen=50, rotindex=True, cache_path=None, raw=False):
        """cache_path: keep / use the parsed arrays in that binary file (vsk_trset_load_leb36_cached);
        `.from_cache` tells whether the text was parsed (False) or the file served (True)"""
        h = C.c_void_p()
        hit = C.c_int(0)
        if raw:  # tokenised only; the GPU decodes profiles and flanks at upload (vsk_trset_load_leb36_raw)
            rc = lib().vsk_trset_load_leb36_raw(path.encode(), int(is_read), maxflank, option_r, reflen, C.byref(h))
        elif cache_path:
            rc = lib().vsk_trset_load_leb36_cached(path.encode(), int(is_read), maxflank, option_r, reflen,
                                                   os.fsencode(cache_path), C.byref(h), C.byref(hit))
        else:
            rc = lib().vsk_trset_load_leb36(path.encode(), int(is_read), maxflank, option_r, reflen, C.byref(h))
        if rc != 0:
            raise VskError(f"cannot load {path} ({rc})")
        ts = cls(h)
        ts.from_cache = bool(hit.value)
        if rotindex:
            ts.apply_rotindex(path + ".rotindex")
        return ts

    @classmethod
    def from_packed(cls, p, is_read, maxflank=50, option_r=0, reflen=50):
        """p: synth.PackedTRSet (or anything with the same arrays)"""
        h = C.c_void_p()
        rc = lib().vsk_trset_from_arrays(len(p.key), p.key, p.patlen, p.prof_off, p.ids, p.flank_off,
                                         p.flanks, int(is_read), maxflank, option_r, reflen, C.byref(h))
        if rc != 0:
            raise VskError(f"vsk_trset_from_arrays failed ({rc})")
        ts = cls(h)
        if getattr(p, "rot", None):
            import os
            import tempfile
            fd, path = tempfile.mkstemp(suffix=".rotindex")
            with os.fdopen(fd, "w") as f:
                for grp in p.rot:
                    f.write(" ".join("%d%s" % (k, d) for k, d in grp) + "\n")
            try:
                ts.apply_rotindex(path)
            finally:
                os.unlink(path)
        return ts

    def apply_rotindex(self, path):
        rc = lib().vsk_trset_apply_rotindex(self.h, path.encode())
        if rc != 0:
            raise VskError(lib().vsk_trset_error(self.h).decode())

    def pin(self):
        """page-lock the host arrays (vsk_trset_pin)"""
        if lib().vsk_trset_pin(self.h) != 0:
            raise VskError("vsk_trset_pin failed")
        return self

    def checksum(self):
        return lib().vsk_trset_checksum(self.h)

    def __len__(self):
        return lib().vsk_trset_size(self.h)

    def close(self):
        if self.h:
            lib().vsk_trset_free(self.h)
            self.h = C.c_void_p()


class RefIndex:
    """vsk_refs: reference TR set + seed index resident in HBM."""

    def __init__(self, ctx, trset, cache_path=None):
        self.ctx, self.trset = ctx, trset
        self.h = C.c_void_p()
        ctx._check(lib().vsk_refs_create_cached(ctx.h, trset.h, os.fsencode(cache_path) if cache_path else None,
                                                C.byref(self.h)))

    @property
    def seed_hits(self):
        return lib().vsk_refs_seed_hits(self.h)

    def close(self):
        if self.h:
            lib().vsk_refs_free(self.ctx.h, self.h)
            self.h = C.c_void_p()


class ResidentReads:
    def __init__(self, ctx, trset):
        self.ctx, self.trset = ctx, trset
        self.h = C.c_void_p()
        ctx._check(lib().vsk_reads_upload(ctx.h, trset.h, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().vsk_reads_free(self.ctx.h, self.h)
            self.h = C.c_void_p()


def _stats_dict(st):
    return {name: int(getattr(st, name)) for name, _ in ProcluStats._fields_}


class _ResultOwner:
    """Frees a vsk_proclu_result when the last array that views its links goes away."""

    def __init__(self, ctx, res):
        self.ctx, self.res = ctx, res

    def __del__(self):
        if self.res and self.ctx.h:  # the links are the context's page-locked memory: not after close()
            lib().vsk_proclu_result_free(self.res)
        self.res = None


class DeviceLinks:
    """The link table of a run, left in HBM (vsk_proclu_result_links_dev): `n` records of
    LINK_DTYPE at device address `ptr`.  It belongs to the context and is overwritten by the
    context's next run.  `.tensor()` views it as a torch int32 tensor [n, 7] without a copy."""

    def __init__(self, ptr, n, device):
        self.ptr, self.n, self.device = ptr, n, device
        nf = LINK_DTYPE.itemsize // 4
        self.__cuda_array_interface__ = {"shape": (n, nf), "typestr": "<i4", "data": (ptr or 0, False), "version": 2}

    def __len__(self):
        return self.n

    def tensor(self):
        import torch
        nf = LINK_DTYPE.itemsize // 4
        if self.n == 0:
            return torch.zeros((0, nf), dtype=torch.int32, device="cuda:%d" % self.device)
        return torch.as_tensor(self, device="cuda:%d" % self.device)


def proclu(ctx, refs, reads, cutoff=88, maxerrors=0, key_budget=0, fetch_links=True, count_cells=False,
           links_on_device=False):
    """vsk_proclu_run (reads = TRSetHandle, host buffers) or vsk_proclu_run_resident
    (reads = ResidentReads).  Returns (links structured array, stats dict).  The array is a
    view of the result's page-locked buffer (no copy); the buffer goes back to the
    context when the array is dropped.  links_on_device=True: the links are not copied to the
    host; a DeviceLinks comes back instead (the send buffer of dist.allgather_links)."""
    prm = ProcluParams(cutoff, maxerrors, 0, int(count_cells), key_budget, int(links_on_device), 0, None, None)
    res = C.c_void_p()
    if isinstance(reads, ResidentReads):
        ctx._check(lib().vsk_proclu_run_resident(ctx.h, refs.h, reads.h, C.byref(prm), C.byref(res)))
    else:
        ctx._check(lib().vsk_proclu_run(ctx.h, refs.h, reads.h, C.byref(prm), C.byref(res)))
    st = ProcluStats()
    lib().vsk_proclu_result_stats(res, C.byref(st))
    links = None
    if links_on_device:
        ptr = C.c_void_p()
        n = lib().vsk_proclu_result_links_dev(res, C.byref(ptr))
        lib().vsk_proclu_result_free(res)
        return DeviceLinks(ptr.value, int(n), ctx.device), _stats_dict(st)
    if fetch_links:
        ptr = C.c_void_p()
        n = lib().vsk_proclu_result_links(res, C.byref(ptr))
        if n:
            buf = (C.c_char * (n * LINK_DTYPE.itemsize)).from_address(ptr.value)
            buf._owner = _ResultOwner(ctx, res)
            links = np.frombuffer(buf, dtype=LINK_DTYPE, count=n)
            return links, _stats_dict(st)
        links = np.zeros(0, LINK_DTYPE)
    lib().vsk_proclu_result_free(res)
    return links, _stats_dict(st)


def index_cache_build(trset, path):
    """vsk_index_cache_build: make sure `path` holds the seed index of this reference set (host only).
    Returns True when the file was already good."""
    hit = C.c_int(0)
    rc = lib().vsk_index_cache_build(trset.h, os.fsencode(path), C.byref(hit))
    if rc != 0:
        raise VskError(f"vsk_index_cache_build failed ({rc})")
    return bool(hit.value)


def index_cache_verify(trset, path):
    """vsk_index_cache_verify: True when the file holds this set's index and equals a fresh build"""
    return lib().vsk_index_cache_verify(trset.h, os.fsencode(path)) == 0


def join_clusters(src, out_path):
    """vsk_join_clusters_dir (src = a folder) or vsk_join_clusters_files (src = a list of .clu
    paths): the reference's join_clusters merge.  Host only, no Context.  Returns the counters."""
    from ._lib import JoinStats
    st = JoinStats()
    err = C.create_string_buffer(600)
    if isinstance(src, (str, bytes, os.PathLike)):
        rc = lib().vsk_join_clusters_dir(os.fsencode(src), os.fsencode(out_path), C.byref(st), 0, err, len(err))
    else:
        arr = (C.c_char_p * len(src))(*[os.fsencode(p) for p in src])
        rc = lib().vsk_join_clusters_files(arr, len(src), os.fsencode(out_path), C.byref(st), err, len(err))
    if rc != 0:
        raise VskError(f"join_clusters failed ({rc}): {err.value.decode()}")
    return {name: int(getattr(st, name)) for name, _ in JoinStats._fields_ if name != "pad_"}


def rankflank(mapdir, rank_tsv, map_tsv=None):
    """vsk_rankflank_dir: best-reference rows (refid, readid, score, ties) from flankalign's .map
    files, and optionally the (refid, readid) rows of every hit.  Host only.  Returns the counters."""
    from ._lib import RankStats
    st = RankStats()
    rc = lib().vsk_rankflank_dir(os.fsencode(mapdir), os.fsencode(rank_tsv),
                                 os.fsencode(map_tsv) if map_tsv else None, C.byref(st))
    if rc != 0:
        raise VskError(f"vsk_rankflank_dir failed ({rc})")
    return {name: int(getattr(st, name)) for name, _ in RankStats._fields_}


def write_clusters(links, clu_path, map_path=None):
    links = np.ascontiguousarray(links, LINK_DTYPE)
    tr = C.c_uint64()
    rc = lib().vsk_clusters_write(links.ctypes.data_as(C.c_void_p), len(links), clu_path.encode(),
                                  map_path.encode() if map_path else None, C.byref(tr))
    if rc != 0:
        raise VskError(f"vsk_clusters_write failed ({rc})")
    return tr.value
