import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import vntrseek_b200 as v
from oracle.pyoracle import Oracle
from vntrseek_b200 import synth
tab = np.load("/root/repo/tests/golden/dist_table.npz")["sm"]
ctx = v.Context(tab, 0); orc = Oracle(tab)
rng = np.random.default_rng(1)
for L in (130, 200, 300, 513, 700, 1100):
    n = 64
    lens = np.full(2*n, L)
    off, ids = synth._random_profiles(rng, lens)
    ids = ids.astype(np.uint16)
    ids[off[1]:off[2]] = ids[off[0]:off[1]]
    li = np.arange(0, 2*n, 2, dtype=np.int32); ti = li + 1
    for mode in (0, 1):
        m = np.full(n, mode, np.uint8)
        try:
            ps = ctx.profiles(off=off, ids=ids)
            got = ctx.pair_dp(ps, ps, li, ti, m, 88)
        except Exception as e:
            print("L", L, "mode", mode, "EXC", str(e)[:150]); sys.exit(0)
        exp = orc.dp_batch(off, ids, off, ids, li, ti, m, 88)
        bad = np.nonzero((got != exp).any(1))[0]
        print("L", L, "mode", mode, "bad", len(bad), got[bad[:2]].tolist(), exp[bad[:2]].tolist(), flush=True)
        ps.close()
