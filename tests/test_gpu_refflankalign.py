"""GPU, file level: the drop-in refflankalign.exe against the reference binary (oracle/_ref):
every `<a>_<b>.xml` and representatives.txt byte for byte; and the banded distance kernel against
the golden vectors recorded from the reference function."""
import filecmp
import os

import numpy as np
import pytest

from _cli import OUR_BIN, REF_BIN, run
from conftest import golden

pytestmark = pytest.mark.gpu


def _blocks(nblocks, seed, flank=30):
    """blocks whose references share (mutated) flanks, so left/right connections occur"""
    from vntrseek_b200 import synth
    rng = np.random.default_rng(seed)
    blocks = []
    for b in range(nblocks):
        base = int(rng.integers(7, 80))
        fl0, fr0 = synth._rand_dna(rng, flank, 0.01), synth._rand_dna(rng, flank, 0.01)
        refs = []
        for r in range(int(rng.integers(1, 12))):
            L = max(2, int(base * rng.uniform(0.8, 1.25)))
            pat = synth._rand_dna(rng, L)
            fl, fr = fl0, fr0
            u = rng.random()
            if u < 0.3:
                fl = synth._rand_dna(rng, flank)
            elif u < 0.6:
                fl = "".join(c if rng.random() > 0.05 else "ACGT"[int(rng.integers(0, 4))] for c in fl0)
                fr = "".join(c if rng.random() > 0.05 else "ACGT"[int(rng.integers(0, 4))] for c in fr0)
            elif u < 0.7:   # opposite strand copy: left = revcomp(right0), right = revcomp(left0)
                fl, fr = synth.revcomp(fr0), synth.revcomp(fl0)
            if rng.random() < 0.2:
                fl = fl.lower()
            refs.append((-(1000 * b + r + 1), "'", fl, pat * 2, fr, pat if rng.random() > 0.2 else pat.lower()))
        reads = [(q + 1, "'", 5, 9, "ACGTACGTACGT", "ACG") for q in range(int(rng.integers(0, 4)))]
        blocks.append((b + 1, 1, refs, reads))
    return blocks


@pytest.mark.parametrize("seed,maxerr,paterr,reflen", [(1, 0, 15, 30), (2, 2, 15, 20), (3, 5, 40, 30)])
def test_refflankalign_matches_reference(tmp_path, seed, maxerr, paterr, reflen):
    if not os.path.exists(os.path.join(REF_BIN, "refflankalign.exe")):
        pytest.skip("oracle/_ref/refflankalign.exe not built")
    from vntrseek_b200 import synth
    inp = str(tmp_path / "in.flanks")
    synth.write_flanks(inp, _blocks(40, seed))
    outs = {}
    for tag, exe in (("ref", os.path.join(REF_BIN, "refflankalign.exe")), ("our", os.path.join(OUR_BIN, "refflankalign.exe"))):
        o, x = str(tmp_path / (tag + "_out")), str(tmp_path / (tag + "_xml"))
        os.makedirs(o)
        os.makedirs(x)
        rc, out, err = run(exe, [o, x, inp, maxerr, 2, paterr, reflen], str(tmp_path))
        assert rc == 0, (tag, out[-400:], err[-400:])
        outs[tag] = (o, x)
    assert filecmp.cmp(os.path.join(outs["ref"][0], "representatives.txt"), os.path.join(outs["our"][0], "representatives.txt"),
                       shallow=False)
    fa, fb = sorted(os.listdir(outs["ref"][1])), sorted(os.listdir(outs["our"][1]))
    assert fa == fb and len(fa) == 40
    for f in fa:
        assert filecmp.cmp(os.path.join(outs["ref"][1], f), os.path.join(outs["our"][1], f), shallow=False), f
    xml = open(os.path.join(outs["our"][1], fa[0])).read()
    assert "<cluster trim=" in xml


def test_nbdist_golden(ctx):
    import ctypes as C

    from vntrseek_b200 import pack_strings
    from vntrseek_b200._lib import lib
    g = golden("kat_nbdist.npz")
    a, b, me, exp = list(g["a"]), list(g["b"]), g["maxerr"], g["expect"]
    L = lib()
    L.vsk_nbdist_pairs.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
    for k in sorted(set(int(x) for x in me)):
        if k > 31:
            continue
        idx = [i for i in range(len(a)) if me[i] == k]
        abuf, aoff = pack_strings([a[i] for i in idx])
        bbuf, boff = pack_strings([b[i] for i in idx])
        out = np.zeros(len(idx), np.int32)
        rc = L.vsk_nbdist_pairs(ctx.h, abuf, aoff.ctypes.data, bbuf, boff.ctypes.data, len(idx), k, out.ctypes.data)
        assert rc == 0
        assert (out == exp[idx]).all(), (k, np.nonzero(out != exp[idx])[0][:5])
