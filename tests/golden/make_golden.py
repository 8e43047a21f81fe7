"""Regenerates the golden fixtures in this directory FROM THE REFERENCE.

Runs only where /root/reference exists (the build container): it loads the
unmodified reference functions through oracle/_ref/libref_harness.so (built by
`make -C oracle ref` from the sources where they lie) and records their outputs
on seeded inputs.  The GPU box has no /root/reference; tests there read the
committed .npz files.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.pyoracle import Reference, build  # noqa: E402
from vntrseek_b200 import synth  # noqa: E402


def rand_dna(rng, n, alpha=b"ACGT"):
    return bytes(rng.choice(list(alpha), n).tolist())


def mutate(rng, s, rate):
    out = bytearray()
    for ch in s:
        u = rng.random()
        if u < rate / 3:
            continue
        if u < 2 * rate / 3:
            out.append(int(rng.choice(list(b"ACGT"))))
            continue
        out.append(ch)
        if u < rate:
            out.append(int(rng.choice(list(b"ACGT"))))
    return bytes(out)


def main():
    build(ref=True)
    R = Reference()
    table = np.fromfile("/root/reference/src/eucledian.dst", dtype=np.uint8)
    assert table[:12].view(np.int32).tolist() == [777, 10, 5]
    np.savez_compressed(os.path.join(HERE, "dist_table.npz"), sm=table[12:])

    # ---- DP known answers -------------------------------------------------
    rng = np.random.default_rng(20261106)
    profs, pairs = [], []

    def add_pair(left, top, modes):
        profs.append(np.asarray(left, np.int32))
        profs.append(np.asarray(top, np.int32))
        for m in modes:
            pairs.append((len(profs) - 2, len(profs) - 1, m))

    lens = list(range(1, 40)) + [63, 64, 65, 127, 128, 129, 255, 256, 257, 300, 511, 513, 700]
    for L in lens:
        bases = rng.integers(0, 4, L)
        ids = synth._column_ids(rng, bases)
        mut = synth._mutate_profile(rng, ids, 0.05, 0.03) if L >= 7 else ids.copy()
        ok_nb = max(len(ids), len(mut)) <= 2 * min(len(ids), len(mut))
        add_pair(ids, mut, (0, 1) if ok_nb else (0,))
        # unrelated, random ids over the whole alphabet (large local alphabets)
        L2 = int(rng.integers(max(1, L // 2), 2 * L + 1))
        add_pair(rng.integers(0, 1001, L), rng.integers(0, 1001, L2), (0, 1))
        # rotated copy (band matters)
        if L >= 8:
            add_pair(ids, np.roll(ids, L // 3), (0, 1))
    # extreme ratios, full only
    add_pair(rng.integers(0, 1001, 5), rng.integers(0, 1001, 400), (0,))
    add_pair(rng.integers(0, 1001, 400), rng.integers(0, 1001, 3), (0,))
    add_pair([0, 0, 0], [0, 0], (0,))  # all-dash columns: sumw == 0
    exp = np.zeros((len(pairs), 4), np.int32)
    sim = np.zeros(len(pairs), np.float64)
    for k, (a, b, m) in enumerate(pairs):
        out, s = R.dp(profs[a], profs[b], m, 88)
        exp[k] = out
        sim[k] = s
    off = np.zeros(len(profs) + 1, np.int64)
    off[1:] = np.cumsum([len(p) for p in profs])
    np.savez_compressed(os.path.join(HERE, "kat_dp.npz"), off=off,
                        ids=np.concatenate(profs).astype(np.uint16),
                        pairs=np.array(pairs, np.int32), expect=exp, similarity=sim)

    # ---- LCS ----------------------------------------------------------------
    a, b, e = [], [], []
    for k in range(3000):
        n = int(rng.integers(1, 200)) if k % 4 else int(rng.integers(56, 70))
        s1 = rand_dna(rng, n, b"ACGTN" if k % 11 == 0 else b"ACGT")
        s2 = mutate(rng, s1, 0.15) if k % 2 else rand_dna(rng, int(rng.integers(1, 200)))
        s2 = s2 or b"A"
        a.append(s1)
        b.append(s2)
        e.append(R.lcs(s1, s2))
    np.savez_compressed(os.path.join(HERE, "kat_lcs.npz"), a=np.array(a, dtype=object),
                        b=np.array(b, dtype=object), expect=np.array(e, np.int32), allow_pickle=True)

    # ---- edit distance, both copies ----------------------------------------------
    t, p, e0, e1 = [], [], [], []
    for k in range(4000):
        n = int(rng.integers(1, 80)) if k % 3 else int(rng.integers(100, 420))
        T = rand_dna(rng, n, b"ACGTN" if k % 7 == 0 else b"ACGT")
        if k % 2:
            P = mutate(rng, T[:int(rng.integers(1, n + 1))], 0.1)
        else:
            P = rand_dna(rng, int(rng.integers(0, min(n, 80) + 1)))
        t.append(T)
        p.append(P)
        e0.append(R.ed(T, P, 0))
        e1.append(R.ed(T, P, 1))
    # an invalid character on either side
    for T, P in ((b"ACGX", b"AC"), (b"ACGT", b"AcG"), (b"acgt", b"")):
        t.append(T)
        p.append(P)
        e0.append(R.ed(T, P, 0))
        e1.append(R.ed(T, P, 1))
    np.savez_compressed(os.path.join(HERE, "kat_ed.npz"), t=np.array(t, dtype=object),
                        p=np.array(p, dtype=object), expect0=np.array(e0, np.int32),
                        expect1=np.array(e1, np.int32), allow_pickle=True)

    # ---- refflankalign banded DP ----------------------------------------------------
    s1s, s2s, me, e = [], [], [], []
    for k in range(1500):
        n = int(rng.integers(1, 70))
        s1 = rand_dna(rng, n)
        s2 = bytearray(s1)
        for _ in range(int(rng.integers(0, 6))):
            s2[int(rng.integers(0, n))] = int(rng.choice(list(b"ACGT")))
        if k % 5 == 0:
            s2 = bytearray(mutate(rng, s1, 0.1) or b"A")
        m = int(rng.integers(0, n + 3))
        s1s.append(s1)
        s2s.append(bytes(s2))
        me.append(m)
        e.append(R.nbdist(s1, bytes(s2), m))
    np.savez_compressed(os.path.join(HERE, "kat_nbdist.npz"), a=np.array(s1s, dtype=object),
                        b=np.array(s2s, dtype=object), maxerr=np.array(me, np.int32),
                        expect=np.array(e, np.int32), allow_pickle=True)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()


def psearch_small():
    """tests/golden/psearch_small: a small synthetic .leb36 pair (+.rotindex) and what the
    reference psearch.exe (oracle/_ref, unmodified sources) made of it: .clu + stderr log."""
    import subprocess
    d = os.path.join(HERE, "psearch_small")
    os.makedirs(d, exist_ok=True)
    refs = synth.make_refs(250, seed=11, max_patlen=300, rot_frac=0.04)
    reads = synth.make_reads(refs, 300, seed=12, rot_frac=0.03)
    refs.write_leb36(os.path.join(d, "reference.leb36"))
    reads.write_leb36(os.path.join(d, "reads.leb36"))
    p = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "psearch.exe"), "reference.leb36", "reads.leb36",
                        os.path.join(ROOT, "oracle", "_ref", "eucledian.dst"), "88", "0", "50"],
                       cwd=d, capture_output=True, text=True, check=True)
    open(os.path.join(d, "ref.log"), "w").write(p.stderr)


if __name__ == "__main__":
    psearch_small()


def flankalign_small():
    """tests/golden/flankalign_small: a small .flanks file and the .map files the reference
    flankalign.exe (oracle/_ref, unmodified sources) wrote for it."""
    import shutil
    import subprocess
    d = os.path.join(HERE, "flankalign_small")
    os.makedirs(d, exist_ok=True)
    synth.write_flanks(os.path.join(d, "in.flanks"), synth.make_flank_blocks(12, seed=9))
    shutil.rmtree(os.path.join(d, "expected"), ignore_errors=True)
    subprocess.run([os.path.join(ROOT, "oracle", "_ref", "flankalign.exe"), os.path.join(d, "expected"), "xml",
                    os.path.join(d, "in.flanks"), "0", "50", "2", "15"], cwd=d, check=True, capture_output=True)


if __name__ == "__main__":
    flankalign_small()
