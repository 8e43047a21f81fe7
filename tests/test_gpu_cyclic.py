"""GPU: the cyclic aligner vsk_pair_dp_cyclic (GetBestProfileAP, profile.h:1586-1675) through the C ABI
against the known answers recorded from the unmodified reference function (tests/golden/kat_cyclic.npz)
and against the CPU restatement on seeded pairs (rotated / mutated copies, unrelated profiles,
whole-alphabet profiles, sizes up to the limit of 1024 columns)."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


def test_cyclic_golden(ctx):
    g = golden("kat_cyclic.npz")
    ps = ctx.profiles(off=g["off"], ids=g["ids"])
    n = len(g["expect"])
    li = np.arange(0, 2 * n, 2, dtype=np.int32)
    got = ctx.pair_dp_cyclic(ps, ps, li, li + 1)
    bad = np.nonzero((got != g["expect"]).any(1))[0]
    assert len(bad) == 0, (bad[:10], got[bad[:3]], g["expect"][bad[:3]])
    ps.close()


def test_cyclic_vs_oracle_random(ctx, oracle):
    from vntrseek_b200 import synth
    rng = np.random.default_rng(31)
    profs = []
    for k in range(400):
        L = int(rng.integers(1, 200)) if k % 20 else int(rng.integers(600, 1025))
        a = synth._column_ids(rng, rng.integers(0, 4, L))
        kind = k % 4
        if kind == 0 and L >= 7:
            b = np.roll(synth._mutate_profile(rng, a, 0.05, 0.03), int(rng.integers(0, L)))
        elif kind == 1:
            b = np.roll(a, int(rng.integers(0, L)))
        elif kind == 2:
            a = rng.integers(0, 1001, L)
            b = rng.integers(0, 1001, int(rng.integers(1, min(2 * L + 2, 1025))))
        else:
            b = synth._column_ids(rng, rng.integers(0, 4, int(rng.integers(1, 300))))
        profs += [np.asarray(a, np.int32)[:1024], np.asarray(b, np.int32)[:1024]]
    ps = ctx.profiles(profs)
    li = np.arange(0, len(profs), 2, dtype=np.int32)
    got = ctx.pair_dp_cyclic(ps, ps, li, li + 1)
    for k in range(len(li)):
        exp = oracle.cyclic(profs[2 * k], profs[2 * k + 1])
        assert tuple(int(x) for x in got[k]) == exp, (k, len(profs[2 * k]), len(profs[2 * k + 1]))
    assert (got[:, 4] != 0).sum() > 100
    ps.close()


def test_cyclic_limits_and_empty(ctx):
    import vntrseek_b200 as v
    ps = ctx.profiles([np.zeros(1025, np.int32) + 5, np.arange(7, dtype=np.int32)])
    with pytest.raises(v.VskError):
        ctx.pair_dp_cyclic(ps, ps, [0], [1])          # 1025 rows: over the limit, refused
    assert ctx.pair_dp_cyclic(ps, ps, [], []).shape == (0, 5)
    got = ctx.pair_dp_cyclic(ps, ps, [1], [1])
    assert tuple(got[0]) == (0, 0, int(got[0][2]), 7, 0)   # a profile against itself: distance 0, rotation 0
    ps.close()
