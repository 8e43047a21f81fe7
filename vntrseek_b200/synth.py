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
