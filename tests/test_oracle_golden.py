"""CPU: the oracle restatement (oracle/vsk_oracle.c) against the golden vectors
recorded from the unmodified reference functions (tests/golden/make_golden.py),
and -- where oracle/_ref exists -- against the reference functions live."""
import numpy as np

from conftest import golden


def _profiles(g):
    off, ids = g["off"], g["ids"].astype(np.int32)
    return [ids[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def test_dp_golden(oracle):
    g = golden("kat_dp.npz")
    profs = _profiles(g)
    for (a, b, m), exp, sim in zip(g["pairs"], g["expect"], g["similarity"]):
        got = oracle.dp(profs[a], profs[b], int(m), 88)
        assert got == tuple(int(x) for x in exp), (a, b, m)
        if exp[2] != 0:
            assert oracle.similarity(got[1], got[2]) == sim  # bit-exact double


def test_lcs_golden(oracle):
    g = golden("kat_lcs.npz")
    for a, b, e in zip(g["a"], g["b"], g["expect"]):
        assert oracle.lcs(a, b) == e, (a, b)


def test_lcs_quirk_61_63(oracle):
    # single-word copy starts its bitmask at 0x10: longer length 61..63 -> only 60 chars seen
    s = b"ACGT" * 16
    for n in (60, 61, 62, 63):
        assert oracle.lcs(s[:n], s[:n]) == min(n, 60)
    assert oracle.lcs(s[:64], s[:64]) == 64


def test_ed_golden(oracle):
    g = golden("kat_ed.npz")
    for t, p, e0, e1 in zip(g["t"], g["p"], g["expect0"], g["expect1"]):
        assert oracle.ed(t, p, 0) == e0, (t, p)
        assert oracle.ed(t, p, 1) == e1, (t, p)


def test_ed_copies_differ_on_long_text():
    g = golden("kat_ed.npz")
    assert (g["expect0"] != g["expect1"]).any()  # the stale-carry copy deviates for N >= 128


def test_nbdist_golden(oracle):
    g = golden("kat_nbdist.npz")
    for a, b, m, e in zip(g["a"], g["b"], g["maxerr"], g["expect"]):
        assert oracle.nbdist(a, b, int(m)) == e, (a, b, m)


def test_thresholds(oracle):
    # hclust.h:537-541 and hclust.h:664-667
    assert oracle.lcs_min(20, 25) == 17
    assert oracle.lcs_min(7, 7) == 6
    assert [oracle.flank_maxerr(n) for n in (0, 10, 19, 20, 50)] == [0, 4, 7, 8, 8]


def test_oracle_vs_reference_live(oracle, reference):
    rng = np.random.default_rng(7)
    for it in range(400):
        l1 = int(rng.integers(1, 150))
        l2 = int(rng.integers(max(1, l1 // 2), 2 * l1 + 1))
        left = rng.integers(0, 1001, l1)
        top = rng.integers(0, 1001, l2)
        for m in (0, 1):
            assert oracle.dp(left, top, m) == reference.dp(left, top, m)[0]


def test_cyclic_golden(oracle):
    """the cyclic aligner's restatement (vo_cyclic) against known answers recorded from the unmodified
    GetBestProfileAP (tests/golden/make_cyclic_golden.py): distance, weighted sums, length, rotation"""
    g = golden("kat_cyclic.npz")
    profs = _profiles(g)
    for k, (exp, sim) in enumerate(zip(g["expect"], g["similarity"])):
        got = oracle.cyclic(profs[2 * k], profs[2 * k + 1])
        assert got == tuple(int(x) for x in exp), (k, len(profs[2 * k]), len(profs[2 * k + 1]))
        if exp[2] != 0:
            assert oracle.similarity(got[1], got[2]) == sim
    assert (g["expect"][:, 4] != 0).sum() > 100


def test_cyclic_vs_reference_live(oracle, reference):
    from vntrseek_b200 import synth
    rng = np.random.default_rng(77)
    for k in range(150):
        L = int(rng.integers(1, 150))
        a = synth._column_ids(rng, rng.integers(0, 4, L)).astype(np.int32)
        b = np.roll(synth._mutate_profile(rng, a, 0.08, 0.04), int(rng.integers(0, L))) if (k % 2 and L >= 7) else \
            rng.integers(0, 1001, int(rng.integers(1, 200)))
        b = np.asarray(b, np.int32)
        assert oracle.cyclic(a, b) == reference.cyclic(a, b)[0], (k, len(a), len(b))
