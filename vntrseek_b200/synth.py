"""Synthetic VNTRseek inputs: hg38-shaped reference TR
sets, read TR sets derived from them, and writers for the reference's own file
formats (.leb36 / .rotindex / .edgein / .flanks), so the very same files feed
both the reference binaries and the new ones.

Format facts restated from the reference (no code copied):
  * standard compositions: all (a,c,g,t) with a+c+g+t <= 10 enumerated in nested
    a,c,g,t order -> ids 0..1000 (src/psearch1.91/profile.h:837-886)
  * .leb36 line: KEY PATLEN COPYNUM PROFLEN PROFLENRC LEB36 LEB36RC nA nC nG nT
    LEFT|RIGHT, two little-endian base-36 digits per column (profile.h:44-66)
"""
import numpy as np

_DIG = "0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ"


def _build_compositions():
    counts = []
    for a in range(11):
        for c in range(11):
            for g in range(11):
                for t in range(11):
                    if a + c + g + t <= 10:
                        counts.append((a, c, g, t))
    counts = np.array(counts, dtype=np.int16)
    lut = {tuple(x): i for i, x in enumerate(counts.tolist())}
    comp = np.array([lut[(t, g, c, a)] for a, c, g, t in counts.tolist()], dtype=np.int32)
    return counts, lut, comp


COUNTS, COMP_LUT, COMPLEMENT_ID = _build_compositions()  # [1001,4], dict, [1001]
NONDASHES = COUNTS.sum(1).astype(np.int32)
# consensus symbol: argmax over A,C,G,T,dash with first-max-wins (profile.h:280-307); 4 = dash
_full = np.concatenate([COUNTS, (10 - COUNTS.sum(1))[:, None]], axis=1)
CONS_SYMBOL = np.argmax(_full, axis=1).astype(np.int8)
CONS_SYMBOL[_full.max(1) == 0] = 0
_LEB = np.array([_DIG[i % 36] + _DIG[i // 36] for i in range(1001)])


def comp_id(a, c, g, t):
    return COMP_LUT[(a, c, g, t)]


def leb36(ids):
    return "".join(_LEB[np.asarray(ids)])


def rc_profile(ids):
    """reverse-complement of a profile (GetReverseComplementProfile semantics:
    reversed order, A<->T and C<->G counts swapped)."""
    return COMPLEMENT_ID[np.asarray(ids)[::-1]]


def consensus(ids):
    s = CONS_SYMBOL[np.asarray(ids)]
    return "".join("ACGT"[k] for k in s[s != 4])


_BASES = np.array(list("ACGT"))


def _rand_dna(rng, n, n_rate=0.0):
    s = rng.integers(0, 4, n)
    out = _BASES[s]
    if n_rate > 0 and n > 0:
        out = out.copy()
        out[rng.random(n) < n_rate] = "N"
    return "".join(out)


def _pattern_length(rng):
    """hg38-shaped mixture: .55 in 7-30, .30 in 31-100, .12 in 101-400, .03 in 401-2000"""
    u = rng.random()
    if u < 0.55:
        return int(rng.integers(7, 31))
    if u < 0.85:
        return int(rng.integers(31, 101))
    if u < 0.97:
        return int(rng.integers(101, 401))
    return int(rng.integers(401, 2001))


def _pattern_bases(rng, L, lowcomplexity):
    if lowcomplexity:
        p = int(rng.integers(2, 7))
        unit = rng.integers(0, 4, p)
        b = np.tile(unit, L // p + 1)[:L].copy()
        m = rng.random(L) < 0.05
        b[m] = rng.integers(0, 4, int(m.sum()))
        return b
    return rng.integers(0, 4, L)


def _column_ids(rng, bases):
    """85% pure columns, 12% two-base mixtures 9/1..7/3, 3% with 1-4 dashes."""
    L = len(bases)
    ids = np.empty(L, np.int32)
    u = rng.random(L)
    for i in range(L):
        cnt = [0, 0, 0, 0]
        b = int(bases[i])
        if u[i] < 0.85:
            cnt[b] = 10
        elif u[i] < 0.97:
            minor = int(rng.integers(1, 4))
            o = (b + int(rng.integers(1, 4))) % 4
            cnt[b] = 10 - minor
            cnt[o] = minor
        else:
            cnt[b] = 10 - int(rng.integers(1, 5))
        ids[i] = COMP_LUT[tuple(cnt)]
    return ids


class TRSet:
    """A set of TR records as a .leb36 file holds them."""

    def __init__(self):
        self.key = []
        self.patlen = []
        self.copynum = []
        self.fwd = []    # list of int32 arrays (profile, forward)
        self.rc = []     # list of int32 arrays (profile, reverse complement)
        self.acgt = []
        self.left = []   # flanks as written in the file (left not reversed)
        self.right = []
        self.rot = []    # rotindex groups: list of [(key, dirchar), ...]

    def __len__(self):
        return len(self.key)

    def add(self, key, ids, left, right, copynum=2.5, patlen=None):
        ids = np.asarray(ids, np.int32)
        self.key.append(int(key))
        cons = CONS_SYMBOL[ids]
        self.patlen.append(int((cons != 4).sum()) if patlen is None else int(patlen))
        self.copynum.append(float(copynum))
        self.fwd.append(ids)
        self.rc.append(rc_profile(ids).astype(np.int32))
        c = consensus(ids)
        self.acgt.append(tuple(c.count(x) for x in "ACGT"))
        self.left.append(left)
        self.right.append(right)

    def write_leb36(self, path):
        with open(path, "w") as f:
            for i in range(len(self)):
                a, c, g, t = self.acgt[i]
                f.write("%d %d %.2f %d %d %s %s %d %d %d %d %s|%s\n" % (
                    self.key[i], self.patlen[i], self.copynum[i], len(self.fwd[i]), len(self.rc[i]),
                    leb36(self.fwd[i]), leb36(self.rc[i]), a, c, g, t, self.left[i], self.right[i]))
        if self.rot:
            with open(path + ".rotindex", "w") as f:
                for grp in self.rot:
                    f.write(" ".join("%d%s" % (k, d) for k, d in grp) + "\n")

    def subset(self, idx):
        s = TRSet()
        for i in idx:
            for name in ("key", "patlen", "copynum", "fwd", "rc", "acgt", "left", "right"):
                getattr(s, name).append(getattr(self, name)[i])
        return s


def make_refs(n, seed=1, rot_frac=0.02, lowcomplexity_frac=0.20, flank=60, max_patlen=2000):
    """Reference TR set: keys -1..-n."""
    rng = np.random.default_rng(seed)
    s = TRSet()
    for i in range(n):
        L = min(_pattern_length(rng), max_patlen)
        bases = _pattern_bases(rng, L, rng.random() < lowcomplexity_frac)
        ids = _column_ids(rng, bases)
        s.add(-(i + 1), ids, _rand_dna(rng, flank, 0.001), _rand_dna(rng, flank, 0.001),
              copynum=float(rng.uniform(1.8, 12.0)))
    # redundancy groups: identical / RC-identical profiles (what `redund.exe -i` collapses)
    n_dup = int(n * rot_frac)
    used = set()
    for _ in range(n_dup):
        m = int(rng.integers(0, n))
        d = int(rng.integers(0, n))
        if m == d or m in used or d in used:
            continue
        used.add(m)
        used.add(d)
        if rng.random() < 0.5:
            s.fwd[d], s.rc[d] = s.fwd[m].copy(), s.rc[m].copy()
            dirc = "'"
        else:
            s.fwd[d], s.rc[d] = s.rc[m].copy(), s.fwd[m].copy()
            dirc = '"'
        s.patlen[d] = s.patlen[m]
        c = consensus(s.fwd[d])
        s.acgt[d] = tuple(c.count(x) for x in "ACGT")
        s.rot.append([(s.key[m], "'"), (s.key[d], dirc)])
    return s


def _mutate_profile(rng, ids, sub=0.03, indel=0.01):
    out = []
    for v in ids:
        u = rng.random()
        if u < indel / 2:
            continue                       # column deleted
        if u < indel:
            out.append(int(v))             # column inserted after
            b = int(rng.integers(0, 4))
            cnt = [0, 0, 0, 0]
            cnt[b] = 10
            out.append(COMP_LUT[tuple(cnt)])
            continue
        if u < indel + sub:
            b = int(rng.integers(0, 4))
            cnt = [0, 0, 0, 0]
            cnt[b] = 10
            out.append(COMP_LUT[tuple(cnt)])
            continue
        out.append(int(v))
    if len(out) < 7:
        out = [int(x) for x in ids]
    return np.array(out, np.int32)


def _mutate_dna(rng, s, sub=0.03, indel=0.005):
    out = []
    for ch in s:
        u = rng.random()
        if u < indel / 2:
            continue
        if u < indel:
            out.append(ch)
            out.append("ACGT"[int(rng.integers(0, 4))])
            continue
        if u < indel + sub:
            out.append("ACGT"[int(rng.integers(0, 4))])
            continue
        out.append(ch)
    return "".join(out)


_CMAP = str.maketrans("ACGTN", "TGCAN")


def revcomp(s):
    return s.translate(_CMAP)[::-1]


def make_reads(refs, n, seed=2, derived_frac=0.70, key0=1, rot_frac=0.01):
    """Read TR set: keys key0..; 70% derived from a random reference TR (same
    rotation; 3% column substitutions, 1% column indels; random orientation;
    flanks = the 10-50 nt of the reference flanks next to the repeat, mutated),
    30% unrelated decoys."""
    rng = np.random.default_rng(seed)
    s = TRSet()
    for i in range(n):
        key = key0 + i
        if rng.random() < derived_frac and len(refs):
            r = int(rng.integers(0, len(refs)))
            ids = _mutate_profile(rng, refs.fwd[r])
            ll = int(rng.integers(10, 51))
            rl = int(rng.integers(10, 51))
            left = _mutate_dna(rng, refs.left[r][-ll:])
            right = _mutate_dna(rng, refs.right[r][:rl])
            if rng.random() < 0.5:   # stored in the opposite orientation
                ids = rc_profile(ids).astype(np.int32)
                left, right = revcomp(right), revcomp(left)
        else:
            L = _pattern_length(rng)
            ids = _column_ids(rng, _pattern_bases(rng, L, rng.random() < 0.2))
            left = _rand_dna(rng, int(rng.integers(10, 51)))
            right = _rand_dna(rng, int(rng.integers(10, 51)))
        s.add(key, ids, left, right, copynum=float(rng.uniform(1.8, 8.0)))
    n_dup = int(n * rot_frac)
    used = set()
    for _ in range(n_dup):
        m = int(rng.integers(0, n))
        d = int(rng.integers(0, n))
        if m == d or m in used or d in used:
            continue
        used.add(m)
        used.add(d)
        if rng.random() < 0.5:
            s.fwd[d], s.rc[d] = s.fwd[m].copy(), s.rc[m].copy()
            dirc = "'"
        else:
            s.fwd[d], s.rc[d] = s.rc[m].copy(), s.fwd[m].copy()
            dirc = '"'
        s.patlen[d] = s.patlen[m]
        c = consensus(s.fwd[d])
        s.acgt[d] = tuple(c.count(x) for x in "ACGT")
        s.rot.append([(s.key[m], "'"), (s.key[d], dirc)])
    return s


# ---------------------------------------------------------------------------
# Vectorised generator for the bench-scale sets (250 K refs, 1 M+ reads).
# Same recipe as make_refs / make_reads, produced as packed arrays.
# ---------------------------------------------------------------------------
_PURE = np.array([COMP_LUT[tuple(10 if k == b else 0 for k in range(4))] for b in range(4)], np.int32)
_MIX = np.zeros((4, 4, 4), np.int32)   # [base][other][minor]
_DASH = np.zeros((4, 5), np.int32)     # [base][ndash]
for _b in range(4):
    for _o in range(4):
        for _m in range(1, 4):
            if _o != _b:
                _c = [0, 0, 0, 0]
                _c[_b] = 10 - _m
                _c[_o] = _m
                _MIX[_b, _o, _m] = COMP_LUT[tuple(_c)]
    for _k in range(1, 5):
        _c = [0, 0, 0, 0]
        _c[_b] = 10 - _k
        _DASH[_b, _k] = COMP_LUT[tuple(_c)]
_ASCII = np.frombuffer(b"ACGT", np.uint8)
_COMP_ASCII = np.arange(256, dtype=np.uint8)
for _x, _y in zip(b"ACGTN", b"TGCAN"):
    _COMP_ASCII[_x] = _y


class PackedTRSet:
    """TR records as flat arrays, in the layout vsk_trset_from_arrays takes."""

    def __init__(self, key, patlen, prof_off, ids, flank_off, flanks, rot=()):
        self.key = np.ascontiguousarray(key, np.int32)
        self.patlen = np.ascontiguousarray(patlen, np.int32)
        self.prof_off = np.ascontiguousarray(prof_off, np.int64)    # [2n+1]
        self.ids = np.ascontiguousarray(ids, np.uint16)
        self.flank_off = np.ascontiguousarray(flank_off, np.int64)  # [2n+1]
        self.flanks = np.ascontiguousarray(flanks, np.uint8)
        self.rot = list(rot)

    def __len__(self):
        return len(self.key)

    def slice(self, a, b):
        po = self.prof_off[2 * a:2 * b + 1]
        fo = self.flank_off[2 * a:2 * b + 1]
        keys = set(self.key[a:b].tolist())
        rot = [g for g in self.rot if all(k in keys for k, _ in g)]
        return PackedTRSet(self.key[a:b], self.patlen[a:b], po - po[0], self.ids[po[0]:po[-1]],
                           fo - fo[0], self.flanks[fo[0]:fo[-1]], rot)

    def write_leb36(self, path):
        fl = self.flanks.tobytes().decode()
        with open(path, "w") as f:
            for i in range(len(self)):
                a, b, c = self.prof_off[2 * i:2 * i + 3]
                fwd, rc = self.ids[a:b], self.ids[b:c]
                x, y, z = self.flank_off[2 * i:2 * i + 3]
                s = CONS_SYMBOL[fwd]
                cnt = np.bincount(s[s != 4], minlength=4)
                f.write("%d %d %.2f %d %d %s %s %d %d %d %d %s|%s\n" % (
                    self.key[i], self.patlen[i], 2.5, b - a, c - b, leb36(fwd), leb36(rc),
                    cnt[0], cnt[1], cnt[2], cnt[3], fl[x:y], fl[y:z]))
        if self.rot:
            with open(path + ".rotindex", "w") as f:
                for grp in self.rot:
                    f.write(" ".join("%d%s" % (k, d) for k, d in grp) + "\n")


def _lengths(rng, n, max_patlen=2000):
    u = rng.random(n)
    L = np.where(u < .55, rng.integers(7, 31, n),
                 np.where(u < .85, rng.integers(31, 101, n),
                          np.where(u < .97, rng.integers(101, 401, n), rng.integers(401, 2001, n))))
    return np.minimum(L, max_patlen).astype(np.int64)


def _seg(L):
    """offsets, owner index per element, position inside the segment"""
    off = np.zeros(len(L) + 1, np.int64)
    np.cumsum(L, out=off[1:])
    owner = np.repeat(np.arange(len(L)), L)
    pos = np.arange(off[-1]) - off[owner]
    return off, owner, pos


def _random_profiles(rng, L, lowfrac=0.20):
    n = len(L)
    off, owner, pos = _seg(L)
    tot = int(off[-1])
    bases = rng.integers(0, 4, tot)
    low = rng.random(n) < lowfrac
    period = rng.integers(2, 7, n)
    units = rng.integers(0, 4, (n, 6))
    lc = low[owner]
    tiled = units[owner, pos % period[owner]]
    keep = rng.random(tot) >= 0.05
    bases = np.where(lc & keep, tiled, bases)
    u = rng.random(tot)
    minor = rng.integers(1, 4, tot)
    other = (bases + rng.integers(1, 4, tot)) % 4
    nd = rng.integers(1, 5, tot)
    ids = np.where(u < .85, _PURE[bases], np.where(u < .97, _MIX[bases, other, minor], _DASH[bases, nd]))
    return off, ids.astype(np.int32)


def _rc_segments(ids, off):
    """reverse-complement every segment"""
    L = np.diff(off)
    owner = np.repeat(np.arange(len(L)), L)
    pos = np.arange(off[-1]) - off[owner]
    src = off[owner + 1] - 1 - pos
    return COMPLEMENT_ID[ids[src]]


def _dna(rng, n, n_rate=0.0):
    s = _ASCII[rng.integers(0, 4, n)]
    if n_rate:
        s = np.where(rng.random(n) < n_rate, np.uint8(ord("N")), s)
    return s


def _mutate_segments(rng, vals, off, sub, indel, newval, protect_min=0):
    """column-wise substitutions / deletions / insertions inside segments; returns (vals, off)"""
    L = np.diff(off)
    owner = np.repeat(np.arange(len(L)), L)
    tot = len(vals)
    u = rng.random(tot)
    cnt = np.ones(tot, np.int64)
    cnt[u < indel / 2] = 0
    cnt[(u >= indel / 2) & (u < indel)] = 2
    newlen = np.bincount(owner, weights=cnt, minlength=len(L)).astype(np.int64)
    if protect_min:
        bad = newlen < protect_min
        if bad.any():
            cnt[bad[owner]] = 1
            newlen = np.bincount(owner, weights=cnt, minlength=len(L)).astype(np.int64)
    out = np.repeat(vals, cnt)
    sub_src = (u >= indel) & (u < indel + sub)
    first = np.cumsum(cnt) - cnt            # output index of each source element's first copy
    s_idx = first[sub_src & (cnt > 0)]
    out[s_idx] = newval(len(s_idx))
    ins = first[cnt == 2] + 1
    out[ins] = newval(len(ins))
    noff = np.zeros(len(L) + 1, np.int64)
    np.cumsum(newlen, out=noff[1:])
    return out, noff


def _gather_segments(vals, off, which):
    """concatenate segments `which` of (vals, off)"""
    L = (off[1:] - off[:-1])[which]
    noff, owner, pos = _seg(L)
    return vals[off[which][owner] + pos], noff


def _assemble(key, fwd, foff, rcv, left, loff, right, roff, rot):
    n = len(key)
    Lf = np.diff(foff)
    prof_len = np.empty(2 * n, np.int64)
    prof_len[0::2] = Lf
    prof_len[1::2] = Lf
    prof_off = np.zeros(2 * n + 1, np.int64)
    np.cumsum(prof_len, out=prof_off[1:])
    ids = np.empty(int(prof_off[-1]), np.uint16)
    _, owner, pos = _seg(Lf)
    ids[prof_off[2 * owner] + pos] = fwd
    ids[prof_off[2 * owner + 1] + pos] = rcv
    fl_len = np.empty(2 * n, np.int64)
    fl_len[0::2] = np.diff(loff)
    fl_len[1::2] = np.diff(roff)
    flank_off = np.zeros(2 * n + 1, np.int64)
    np.cumsum(fl_len, out=flank_off[1:])
    flanks = np.empty(int(flank_off[-1]), np.uint8)
    _, o1, p1 = _seg(np.diff(loff))
    flanks[flank_off[2 * o1] + p1] = left
    _, o2, p2 = _seg(np.diff(roff))
    flanks[flank_off[2 * o2 + 1] + p2] = right
    return PackedTRSet(key, Lf.astype(np.int32), prof_off, ids, flank_off, flanks, rot)


def fast_refs(n, seed=1, rot_frac=0.02, flank=60, max_patlen=2000):
    """reference TR set, keys -1..-n (vectorised make_refs)"""
    rng = np.random.default_rng(seed)
    L = _lengths(rng, n, max_patlen)
    off, ids = _random_profiles(rng, L)
    rcv = _rc_segments(ids, off)
    # redundancy groups: a duplicate takes its master's profile (or its reverse complement)
    n_dup = int(n * rot_frac)
    perm = rng.permutation(n)[:2 * n_dup]
    masters, dups = perm[:n_dup], perm[n_dup:]
    asrc = np.arange(n)
    asrc[dups] = masters
    flip = np.zeros(n, bool)
    flip[dups] = rng.random(n_dup) < 0.5
    f1, o1 = _gather_segments(ids, off, asrc)
    r1, _ = _gather_segments(rcv, off, asrc)
    fo, ow, _ = _seg(np.diff(o1))
    fl = flip[ow]
    fwd = np.where(fl, r1, f1)
    rc = np.where(fl, f1, r1)
    key = -(np.arange(n) + 1)
    rot = [[(int(key[m]), "'"), (int(key[d]), '"' if flip[d] else "'")] for m, d in zip(masters, dups)]
    loff = np.arange(n + 1, dtype=np.int64) * flank
    left = _dna(rng, n * flank, 0.001)
    right = _dna(rng, n * flank, 0.001)
    return _assemble(key, fwd, o1, rc, left, loff, right, loff.copy(), rot)


def fast_reads(refs, n, seed=2, derived_frac=0.70, key0=1, rot_frac=0.01):
    """read TR set derived from `refs` (vectorised make_reads)"""
    rng = np.random.default_rng(seed)
    nref = len(refs)
    derived = rng.random(n) < derived_frac
    src = rng.integers(0, nref, n)
    # forward profiles of the sources / fresh decoys
    rfo = refs.prof_off[0::2]
    rlen = (refs.prof_off[1::2] - refs.prof_off[0:-1:2])
    Ld = rlen[src]
    Lr = _lengths(rng, n)
    L = np.where(derived, Ld, Lr)
    off, owner, pos = _seg(L)
    doff, dids = _random_profiles(rng, L)
    srcpos = np.minimum(rfo[src][owner] + pos, len(refs.ids) - 1)
    ids = np.where(derived[owner], refs.ids[srcpos].astype(np.int32), dids)
    # mutate derived reads: 3 % substitutions, 1 % indels (never below 7 columns)
    mids, moff = _mutate_segments(rng, ids, off, 0.03, 0.01, lambda k: _PURE[rng.integers(0, 4, k)], protect_min=7)
    keepd = ~derived
    if keepd.any():  # decoys stay as generated
        pass
    # (decoys were mutated too, which is harmless: they are random anyway)
    rcv = _rc_segments(mids, moff)
    # flanks
    ll = rng.integers(10, 51, n)
    rl = rng.integers(10, 51, n)
    lo, lown, lpos = _seg(ll)
    ro, rown, rpos = _seg(rl)
    rfl = refs.flank_off
    lsrc = rfl[2 * src + 1][lown] - ll[lown] + lpos          # last ll chars of the ref left flank
    rsrc = rfl[2 * src + 1][rown] + rpos                     # first rl chars of the ref right flank
    rand_l = _dna(rng, int(lo[-1]))
    rand_r = _dna(rng, int(ro[-1]))
    left = np.where(derived[lown], refs.flanks[np.clip(lsrc, 0, len(refs.flanks) - 1)], rand_l)
    right = np.where(derived[rown], refs.flanks[np.clip(rsrc, 0, len(refs.flanks) - 1)], rand_r)
    left, lo = _mutate_segments(rng, left, lo, 0.03, 0.005, lambda k: _ASCII[rng.integers(0, 4, k)], protect_min=5)
    right, ro = _mutate_segments(rng, right, ro, 0.03, 0.005, lambda k: _ASCII[rng.integers(0, 4, k)], protect_min=5)
    # half of the reads are stored in the opposite orientation
    flip = rng.random(n) < 0.5
    _, ow, _ = _seg(np.diff(moff))
    fwd = np.where(flip[ow], rcv, mids)
    rc = np.where(flip[ow], mids, rcv)
    # revcomp(right) becomes the left flank and vice versa
    def _revcomp_segments(v, o):
        Lx = np.diff(o)
        own = np.repeat(np.arange(len(Lx)), Lx)
        p = np.arange(o[-1]) - o[own]
        return _COMP_ASCII[v[o[own + 1] - 1 - p]]
    rleft, rright = _revcomp_segments(right, ro), _revcomp_segments(left, lo)
    # assemble per read: choose (left,right) or (revcomp(right), revcomp(left))
    Ll = np.where(flip, np.diff(ro), np.diff(lo))
    Lr2 = np.where(flip, np.diff(lo), np.diff(ro))
    nlo, lown2, lpos2 = _seg(Ll)
    nro, rown2, rpos2 = _seg(Lr2)
    newleft = np.where(flip[lown2], rleft[np.minimum(ro[lown2] + lpos2, len(rleft) - 1)],
                       left[np.minimum(lo[lown2] + lpos2, len(left) - 1)])
    newright = np.where(flip[rown2], rright[np.minimum(lo[rown2] + rpos2, len(rright) - 1)],
                        right[np.minimum(ro[rown2] + rpos2, len(right) - 1)])
    key = key0 + np.arange(n)
    # redundancy groups among the reads
    n_dup = int(n * rot_frac)
    perm = rng.permutation(n)[:2 * n_dup]
    masters, dups = perm[:n_dup], perm[n_dup:]
    asrc = np.arange(n)
    asrc[dups] = masters
    dflip = np.zeros(n, bool)
    dflip[dups] = rng.random(n_dup) < 0.5
    f1, o1 = _gather_segments(fwd, moff, asrc)
    r1, _ = _gather_segments(rc, moff, asrc)
    _, ow2, _ = _seg(np.diff(o1))
    ff = np.where(dflip[ow2], r1, f1)
    rr = np.where(dflip[ow2], f1, r1)
    rot = [[(int(key[m]), "'"), (int(key[d]), '"' if dflip[d] else "'")] for m, d in zip(masters, dups)]
    return _assemble(key, ff, o1, rr, newleft, nlo, newright, nro, rot)


# ---------------------------------------------------------------------------
# .flanks blocks (the input of flankalign.exe, written by perl/run_flankcomp.pl:229-268)
# ---------------------------------------------------------------------------
def make_flank_blocks(nblocks, nref=6, nread=40, seed=5, mixed_case=True):
    """list of (cluster, part, refs, reads); refs = (rid, dir, FL, SEQ, FR, PAT),
    reads = (rid, dir, first, last, DNA, PAT).  Reads carry 10-70 nt of a reference's
    flanks (mutated), half of them reverse-complemented; pattern sizes inside a block
    spread enough that the +-15 % window matters."""
    rng = np.random.default_rng(seed)
    blocks = []
    rid_read = 1
    for b in range(nblocks):
        base = int(rng.integers(7, 120))
        refs = []
        for r in range(int(rng.integers(1, nref + 1))):
            L = max(2, int(base * rng.uniform(0.7, 1.4)))
            pat = _rand_dna(rng, L)
            fl, fr = _rand_dna(rng, int(rng.integers(20, 61)), 0.01), _rand_dna(rng, int(rng.integers(20, 61)), 0.01)
            if mixed_case and rng.random() < 0.3:
                fl = fl.lower()
            refs.append((-(1000 * b + r + 1), "'", fl, pat * 2, fr, pat))
        reads = []
        for q in range(int(rng.integers(1, nread + 1))):
            src = refs[int(rng.integers(0, len(refs)))]
            pat = src[5]
            if rng.random() < 0.5:
                L = max(2, int(len(pat) * rng.uniform(0.8, 1.25)))
                pat = _rand_dna(rng, L)
            ll, rl = int(rng.integers(0, 71)), int(rng.integers(0, 71))
            left = _mutate_dna(rng, src[2].upper()[-ll:] if ll else "", 0.04, 0.01)
            right = _mutate_dna(rng, src[4].upper()[:rl], 0.04, 0.01)
            if rng.random() < 0.3:
                left = _rand_dna(rng, len(left))
            rep = pat * int(rng.integers(2, 5))
            dna = left + rep + right
            first, last = len(left) + 1, len(left) + len(rep)
            if rng.random() < 0.5:
                dna = revcomp(dna.replace("N", "A"))
                first, last = len(right) + 1, len(right) + len(rep)
            reads.append((rid_read, '"' if rng.random() < 0.5 else "'", first, last, dna, pat))
            rid_read += 1
        blocks.append((b + 1, 1, refs, reads))
    return blocks


def write_flanks(path, blocks):
    with open(path, "w") as f:
        for c, h, refs, reads in blocks:
            f.write("@(%d_%d):\n" % (c, h))
            f.write("*" * 70 + "\n")
            for rid, d, fl, seq, fr, pat in refs:
                f.write("%d%s,%s,%s,%s,%s\n" % (rid, d, fl, seq, fr, pat))
            for rid, d, first, last, dna, pat in reads:
                f.write("%d%s,%d,%d,%s,%s\n" % (rid, d, first, last, dna, pat))
