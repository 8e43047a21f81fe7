"""Makes tests/golden/kat_cyclic.npz FROM THE REFERENCE: known answers of the cyclic aligner
GetBestProfileAP (src/psearch1.91/profile.h:1586-1675), called through oracle/_ref/libref_harness.so
(the unmodified function, compiled where it lies) on seeded pairs: rotated copies, rotated + mutated
copies, unrelated profiles, whole-alphabet profiles, lengths 1..400 incl. the kernel's class limits.

    python tests/golden/make_cyclic_golden.py        (needs /root/reference)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.pyoracle import Reference, build  # noqa: E402
from vntrseek_b200 import synth  # noqa: E402


def main():
    build(ref=True)
    R = Reference()
    rng = np.random.default_rng(20261107)
    profs, expect, sims = [], [], []
    lens = list(range(1, 34)) + [47, 63, 64, 65, 100, 127, 128, 129, 200, 255, 256, 257, 300, 400]
    for L in lens:
        for kind in range(5):
            a = synth._column_ids(rng, rng.integers(0, 4, L)).astype(np.int32)
            if kind == 0:
                b = np.roll(a, int(rng.integers(0, L)))
            elif kind == 1 and L >= 7:
                b = np.roll(synth._mutate_profile(rng, a, 0.05, 0.03), int(rng.integers(0, L)))
            elif kind == 2:
                a = rng.integers(0, 1001, L).astype(np.int32)
                b = rng.integers(0, 1001, int(rng.integers(1, min(2 * L + 2, 401))))
            elif kind == 3:
                b = synth._column_ids(rng, rng.integers(0, 4, int(rng.integers(1, 200))))
            else:
                b = synth.rc_profile(np.roll(a, int(rng.integers(0, L))))
            b = np.asarray(b, np.int32)
            out, sim = R.cyclic(a, b)
            profs += [a, b]
            expect.append(out)
            sims.append(sim)
    off = np.zeros(len(profs) + 1, np.int64)
    np.cumsum([len(p) for p in profs], out=off[1:])
    np.savez_compressed(os.path.join(HERE, "kat_cyclic.npz"), off=off, ids=np.concatenate(profs).astype(np.uint16),
                        expect=np.array(expect, np.int32), similarity=np.array(sims, np.float64))
    e = np.array(expect)
    print("pairs", len(expect), "with rotation != 0:", int((e[:, 4] != 0).sum()))


if __name__ == "__main__":
    main()
