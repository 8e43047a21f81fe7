"""GPU parity: the CUDA kernels, called through the C ABI, against the golden
vectors recorded from the reference and against the CPU oracle on seeded inputs."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


def _profiles(g):
    off, ids = g["off"], g["ids"].astype(np.int32)
    return [ids[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def test_dp_golden(ctx):
    g = golden("kat_dp.npz")
    ps = ctx.profiles(off=g["off"], ids=g["ids"])
    pairs = g["pairs"]
    got = ctx.pair_dp(ps, ps, pairs[:, 0], pairs[:, 1], pairs[:, 2].astype(np.uint8), 88)
    bad = np.nonzero((got != g["expect"]).any(1))[0]
    assert len(bad) == 0, (bad[:10], got[bad[:3]], g["expect"][bad[:3]])
    ps.close()


@pytest.mark.parametrize("seed,maxlen", [(1, 60), (2, 300), (3, 1200)])
def test_dp_vs_oracle_random(ctx, oracle, seed, maxlen):
    from vntrseek_b200 import synth
    rng = np.random.default_rng(seed)
    profs, li, ti, mode = [], [], [], []
    n = 300 if maxlen <= 300 else 60
    for k in range(n):
        L = int(rng.integers(1, maxlen + 1))
        ids = synth._column_ids(rng, rng.integers(0, 4, L))
        kind = k % 4
        if kind == 0 and L >= 7:
            other = synth._mutate_profile(rng, ids, 0.05, 0.03)
        elif kind == 1:
            other = np.roll(ids, int(rng.integers(0, L)))
        elif kind == 2:
            other = rng.integers(0, 1001, int(rng.integers(max(1, L // 2), 2 * L + 1)))
            ids = rng.integers(0, 1001, L)
        else:
            other = synth._column_ids(rng, rng.integers(0, 4, int(rng.integers(1, maxlen + 1))))
        profs += [ids, other]
        nb_ok = max(len(ids), len(other)) <= 2 * min(len(ids), len(other))
        for m in ((0, 1) if nb_ok else (0,)):
            li.append(len(profs) - 2)
            ti.append(len(profs) - 1)
            mode.append(m)
    ps = ctx.profiles(profs)
    got = ctx.pair_dp(ps, ps, li, ti, mode, 88)
    for k in range(len(li)):
        exp = oracle.dp(profs[li[k]], profs[ti[k]], mode[k], 88)
        assert tuple(int(x) for x in got[k]) == exp, (k, len(profs[li[k]]), len(profs[ti[k]]), mode[k])
    ps.close()


def test_dp_max_size_and_idempotence(ctx, oracle):
    """longest accepted profiles (many column strips, boundary scratch), run twice."""
    from vntrseek_b200 import synth
    rng = np.random.default_rng(11)
    a = synth._column_ids(rng, rng.integers(0, 4, 5000))
    b = synth._mutate_profile(rng, a, 0.04, 0.02)
    c = rng.integers(0, 1001, 2600)   # big alphabets: global-table path
    d = rng.integers(0, 1001, 1900)
    ps = ctx.profiles([a, b, c, d])
    li, ti, mode = [0, 0, 2, 2], [1, 1, 3, 3], [0, 1, 0, 1]
    got1 = ctx.pair_dp(ps, ps, li, ti, mode, 88)
    got2 = ctx.pair_dp(ps, ps, li, ti, mode, 88)
    assert (got1 == got2).all()
    profs = [a, b, c, d]
    for k in range(4):
        assert tuple(int(x) for x in got1[k]) == oracle.dp(profs[li[k]], profs[ti[k]], mode[k], 88)
    ps.close()


def test_dp_path_leaves_band(ctx, oracle):
    """narrowband pairs whose unrestricted best path runs through the corners the band cuts off:
    the direction-bit kernels must hand them to the cell-masking kernel (same answer as the oracle)"""
    from vntrseek_b200 import synth
    # Y: a nearly empty column (cheap to skip), X: a full one -- the best alignment skips the Y blocks
    X, Y, Z = synth.comp_id(10, 0, 0, 0), synth.comp_id(0, 1, 0, 0), synth.comp_id(0, 0, 5, 5)
    rng = np.random.default_rng(5)
    profs, li, ti = [], [], []
    for n, m in [(400, 200), (200, 100), (900, 500), (100, 60), (1500, 700), (300, 300)]:
        left = np.array([Y] * n + [X] * m)
        top = np.array([X] * m + [Y] * n)
        profs += [left, top]
        li.append(len(profs) - 2)
        ti.append(len(profs) - 1)
        # the same with noise, and a block that only fits far off the diagonal
        l2 = left.copy()
        l2[rng.integers(0, len(l2), len(l2) // 20)] = Z
        t2 = np.concatenate([top[: m // 2], np.array([Z] * (n // 3)), top[m // 2:]])[: 2 * len(l2) - 1]
        profs += [l2, t2]
        li.append(len(profs) - 2)
        ti.append(len(profs) - 1)
    ps = ctx.profiles(profs)
    got = ctx.pair_dp(ps, ps, li, ti, [1] * len(li), 88)
    full = ctx.pair_dp(ps, ps, li, ti, [0] * len(li), 88)
    differ = 0
    for k in range(len(li)):
        exp = oracle.dp(profs[li[k]], profs[ti[k]], 1, 88)
        assert tuple(int(x) for x in got[k]) == exp, (k, len(profs[li[k]]), len(profs[ti[k]]))
        differ += int(tuple(got[k]) != tuple(full[k]))
    assert differ > 0  # the band did change some answers, i.e. the hand-back path ran
    ps.close()


def test_dp_empty_batch(ctx):
    ps = ctx.profiles([[1, 2, 3]])
    assert ctx.pair_dp(ps, ps, [], [], [], 88).shape == (0, 4)
    ps.close()


def test_dp_rejects_bad_input(ctx):
    import vntrseek_b200 as v
    with pytest.raises(v.VskError):
        ctx.profiles([[1, 2, 1001]])
    ps = ctx.profiles([[1, 2, 3]])
    with pytest.raises(v.VskError):
        ctx.pair_dp(ps, ps, [0], [5], [0], 88)
    ps.close()


def test_lcs_golden(ctx):
    g = golden("kat_lcs.npz")
    got = ctx.lcs_pairs(list(g["a"]), list(g["b"]))
    assert (got == g["expect"]).all(), np.nonzero(got != g["expect"])[0][:10]


def test_lcs_long_vs_oracle(ctx, oracle):
    rng = np.random.default_rng(5)
    a, b = [], []
    for k in range(200):
        n = int(rng.integers(1, 3000))
        s = bytes(rng.choice(list(b"ACGT"), n).tolist())
        t = bytearray(s)
        for _ in range(n // 10):
            t[int(rng.integers(0, n))] = int(rng.choice(list(b"ACGT")))
        if k % 3 == 0:
            t = t[: max(1, n - int(rng.integers(0, n)))]
        a.append(s)
        b.append(bytes(t))
    got = ctx.lcs_pairs(a, b)
    exp = np.array([oracle.lcs(x, y) for x, y in zip(a, b)])
    assert (got == exp).all()


def test_ed_golden(ctx):
    g = golden("kat_ed.npz")
    t, p = list(g["t"]), list(g["p"])
    assert (ctx.ed_pairs(t, p, 0) == g["expect0"]).all()
    assert (ctx.ed_pairs(t, p, 1) == g["expect1"]).all()


def test_ed_prefix_lengths(ctx, oracle):
    """psearch passes N and M separately from the buffers (hclust.h:674-683)."""
    rng = np.random.default_rng(9)
    t, p, tl, pl = [], [], [], []
    for k in range(500):
        T = bytes(rng.choice(list(b"ACGTN"), int(rng.integers(1, 70))).tolist())
        P = bytes(rng.choice(list(b"ACGT"), int(rng.integers(1, 60))).tolist())
        t.append(T)
        p.append(P)
        tl.append(int(rng.integers(1, len(T) + 1)))
        pl.append(int(rng.integers(0, len(P) + 1)))
    for v in (0, 1):
        got = ctx.ed_pairs(t, p, v, tl, pl)
        exp = np.array([oracle.ed(a, b, v, n, m) for a, b, n, m in zip(t, p, tl, pl)])
        assert (got == exp).all()
