"""GPU parity at volume: at least 10^6 seeded random pairs per kernel,
through the C ABI, against the batched CPU oracle (oracle/vsk_oracle.c: vo_*_batch).  The
lengths crowd the places where the kernels change code path: 32/33, 64/65 and 128/129 columns
(thread / 8-lane / warp per pair), 256/384/512 columns (strip widths), 58-66 characters (the
single-word LCS with its 0x10 quirk), text length 128 (where the two edit-distance copies part),
length ratios up to the narrowband limit of 2, and flanks that contain N."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _segments(L):
    off = np.zeros(len(L) + 1, np.int64)
    np.cumsum(L, out=off[1:])
    return off


def _dp_pairs(seed, n_small, n_mid, n_big, wide=True):
    """left/top profile sets + pair list: mutated copies, ratio extremes, unrelated pairs"""
    from vntrseek_b200 import synth
    rng = np.random.default_rng(seed)
    edges = np.array([1, 2, 7, 20, 31, 32, 33, 34, 63, 64, 65, 66, 127, 128, 129, 130])
    Ls = np.where(rng.random(n_small) < 0.3, rng.choice(edges, n_small), rng.integers(1, 140, n_small))
    Lm = np.where(rng.random(n_mid) < 0.5, rng.choice([255, 256, 257, 383, 384, 385, 511, 512, 513, 640], n_mid),
                  rng.integers(129, 700, n_mid))
    Lb = rng.integers(700, 2600, n_big)
    L = np.concatenate([Ls, Lm, Lb]).astype(np.int64)
    n = len(L)
    loff, lids = synth._random_profiles(rng, L)
    lids = lids.astype(np.uint16)
    if wide:   # a few columns from the whole alphabet: more than 200 distinct compositions in the data set
        anyid = rng.random(len(lids)) < 0.02
        lids[anyid] = rng.integers(0, 1001, int(anyid.sum()))
    # top profiles: 55 % mutated copies, 25 % random of a length in [L/2, 2L], 20 % unrelated lengths
    kind = rng.random(n)
    mut, moff = synth._mutate_segments(rng, lids.astype(np.int32), loff, 0.05, 0.03,
                                       lambda k: synth._PURE[rng.integers(0, 4, k)], protect_min=1)
    Lt = np.diff(moff)
    lo = np.maximum(1, (L + 1) // 2)
    Lr = rng.integers(lo, 2 * L + 1)
    Lr = np.where(rng.random(n) < 0.2, np.where(rng.random(n) < 0.5, lo, 2 * L), Lr)   # exactly 1:2 and 2:1
    Lu = np.minimum(rng.integers(1, 200, n), 4 * L + 5)
    Ltop = np.where(kind < 0.55, Lt, np.where(kind < 0.80, Lr, Lu)).astype(np.int64)
    toff = _segments(Ltop)
    _, rnd = synth._random_profiles(rng, Ltop)
    tids = rnd.astype(np.uint16)
    owner = np.repeat(np.arange(n), Ltop)
    pos = np.arange(toff[-1]) - toff[owner]
    use_mut = (kind < 0.55)[owner]
    tids[use_mut] = mut[(moff[owner] + pos)[use_mut]].astype(np.uint16)
    li = np.arange(n, dtype=np.int32)
    ti = np.arange(n, dtype=np.int32)
    nb_ok = np.maximum(L, Ltop) <= 2 * np.minimum(L, Ltop)
    # full matrix for everything, narrowband in addition where the reference would use it
    li2 = np.concatenate([li, li[nb_ok]])
    ti2 = np.concatenate([ti, ti[nb_ok]])
    mode = np.concatenate([np.zeros(n, np.uint8), np.ones(int(nb_ok.sum()), np.uint8)])
    return (loff, lids), (toff, tids), li2, ti2, mode


@pytest.mark.parametrize("wide", [False, True])
def test_dp_million_pairs_vs_oracle(ctx, oracle, wide):
    """wide=False: the data set uses 56 compositions, the small-pair kernels take their costs from the dense
    K x K table in shared memory; wide=True: all 1001 occur, they read the global table"""
    (loff, lids), (toff, tids), li, ti, mode = _dp_pairs(101 + int(wide), 600000, 3000, 120, wide)
    assert (len(np.unique(lids)) > 200) == wide
    assert len(li) >= 1000000
    left = ctx.profiles(off=loff, ids=lids)
    top = ctx.profiles(off=toff, ids=tids)
    got = ctx.pair_dp(left, top, li, ti, mode, 88)
    exp = oracle.dp_batch(loff, lids, toff, tids, li, ti, mode, 88)
    bad = np.nonzero((got != exp).any(1))[0]
    assert len(bad) == 0, (len(bad), bad[:5], got[bad[:3]], exp[bad[:3]],
                           [(int(loff[li[k] + 1] - loff[li[k]]), int(toff[ti[k] + 1] - toff[ti[k]]), int(mode[k]))
                            for k in bad[:5]])
    assert int((mode == 1).sum()) > 300000
    left.close()
    top.close()


def _strings(rng, L, alphabet=b"ACGT", n_rate=0.0):
    off = _segments(L)
    buf = np.frombuffer(alphabet, np.uint8)[rng.integers(0, len(alphabet), int(off[-1]))].copy()
    if n_rate:
        buf[rng.random(len(buf)) < n_rate] = ord("N")
    return buf, off


def _mutated_strings(rng, buf, off, rate):
    out = buf.copy()
    hit = rng.random(len(out)) < rate
    out[hit] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(hit.sum()))]
    return out


def test_lcs_million_pairs_vs_oracle(ctx, oracle):
    from vntrseek_b200._lib import lib
    rng = np.random.default_rng(202)
    n = 1000000
    La = np.where(rng.random(n) < 0.4, rng.integers(56, 68, n), rng.integers(1, 260, n))
    La[:1500] = rng.integers(1000, 5000, 1500)      # many 64-column blocks (systolic kernels, several passes)
    La[1500:1600] = rng.integers(4000, 6000, 100)
    a, aoff = _strings(rng, La, b"ACGTN", 0.0)
    # b = a mutated, cut or extended, so that lengths straddle 60..64 independently of a's
    b_full = _mutated_strings(rng, a, aoff, 0.12)
    Lb = np.clip(La + rng.integers(-6, 7, n), 1, None)
    Lb[La > 999] = np.maximum(1, La[La > 999] - rng.integers(0, 400, int((La > 999).sum())))
    boff = _segments(Lb)
    owner = np.repeat(np.arange(n), Lb)
    pos = np.arange(boff[-1]) - boff[owner]
    src = aoff[owner] + np.minimum(pos, La[owner] - 1)
    b = b_full[src]
    ext = pos >= La[owner]
    b[ext] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(ext.sum()))]
    ia = np.arange(n, dtype=np.int32)
    got = np.zeros(n, np.int32)
    ctx._check(lib().vsk_lcs_pairs(ctx.h, a.tobytes(), aoff, b.tobytes(), boff, n, ia, ia, got))
    exp = oracle.lcs_batch(a.tobytes(), aoff, b.tobytes(), boff, ia, ia)
    bad = np.nonzero(got != exp)[0]
    assert len(bad) == 0, (len(bad), bad[:5], got[bad[:5]], exp[bad[:5]], La[bad[:5]], Lb[bad[:5]])
    for L in range(58, 67):   # every length around the single-word limit occurred on both sides
        assert (La == L).any() and (Lb == L).any()


@pytest.mark.parametrize("variant", [0, 1])
def test_ed_million_pairs_vs_oracle(ctx, oracle, variant):
    """both edit-distance copies (psearch: stale carry from 128 text characters on; newflankalign:
    fixed), prefix lengths passed separately from the buffers, N in the text"""
    from vntrseek_b200._lib import lib
    rng = np.random.default_rng(303 + variant)
    n = 1000000
    Lt = np.where(rng.random(n) < 0.85, rng.integers(1, 72, n), rng.integers(100, 330, n))
    Lt[:2000] = rng.integers(126, 131, 2000)
    t, toff = _strings(rng, Lt, b"ACGT", 0.01)
    # pattern = a mutated stretch of the text (so that small distances occur), random length
    Lp = np.minimum(rng.integers(1, 62, n), Lt + 5)
    Lp[Lt >= 100] = rng.integers(1, 200, int((Lt >= 100).sum()))
    poff = _segments(Lp)
    owner = np.repeat(np.arange(n), Lp)
    pos = np.arange(poff[-1]) - poff[owner]
    shift = rng.integers(0, 4, n)
    src = toff[owner] + np.minimum(pos + shift[owner], Lt[owner] - 1)
    p = _mutated_strings(rng, t, toff, 0.08)[src]
    p[p == ord("N")] = ord("A")
    tlen = np.maximum(1, Lt - rng.integers(0, 3, n) * (rng.random(n) < 0.3)).astype(np.int32)
    plen = np.maximum(0, Lp - rng.integers(0, 3, n) * (rng.random(n) < 0.3)).astype(np.int32)
    got = np.zeros(n, np.int32)
    ctx._check(lib().vsk_ed_pairs(ctx.h, t.tobytes(), np.ascontiguousarray(toff[:-1]), tlen, p.tobytes(),
                                  np.ascontiguousarray(poff[:-1]), plen, n, variant, got))
    exp = oracle.ed_batch(t.tobytes(), toff[:-1], tlen, p.tobytes(), poff[:-1], plen, variant)
    bad = np.nonzero(got != exp)[0]
    assert len(bad) == 0, (len(bad), bad[:5], got[bad[:5]], exp[bad[:5]], tlen[bad[:5]], plen[bad[:5]])
    assert (exp >= 0).sum() > 900000 and (exp == 0).any()


def test_nbdist_million_pairs_vs_oracle(ctx, oracle):
    """refflankalign's banded unit-cost DP (equal lengths; -1 otherwise), maxerr 0..31"""
    from vntrseek_b200._lib import lib
    L_ = lib()
    L_.vsk_nbdist_pairs.argtypes = [C.c_void_p, C.c_char_p, np.ctypeslib.ndpointer(np.int64), C.c_char_p,
                                    np.ctypeslib.ndpointer(np.int64), C.c_int64, C.c_int,
                                    np.ctypeslib.ndpointer(np.int32)]
    rng = np.random.default_rng(404)
    total = 0
    for maxerr in (0, 1, 2, 5, 8, 15, 31):
        n = 150000
        La = rng.integers(1, 90, n)
        a, aoff = _strings(rng, La, b"ACGTN")
        b = _mutated_strings(rng, a, aoff, 0.10)
        Lb = La.copy()
        odd = rng.random(n) < 0.02          # unequal lengths: the reference returns -1
        Lb[odd] = np.maximum(1, La[odd] - 1)
        boff = _segments(Lb)
        owner = np.repeat(np.arange(n), Lb)
        pos = np.arange(boff[-1]) - boff[owner]
        b2 = b[aoff[owner] + pos]
        # a few pairs with an indel (shifted copy) so that the band matters
        got = np.zeros(n, np.int32)
        ctx._check(L_.vsk_nbdist_pairs(ctx.h, a.tobytes(), aoff, b2.tobytes(), boff, n, maxerr, got))
        exp = oracle.nbdist_batch(a.tobytes(), aoff, b2.tobytes(), boff, maxerr)
        bad = np.nonzero(got != exp)[0]
        assert len(bad) == 0, (maxerr, len(bad), bad[:5], got[bad[:5]], exp[bad[:5]])
        assert (exp == -1).any()
        total += n
    assert total >= 1000000
