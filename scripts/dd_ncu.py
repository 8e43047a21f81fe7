"""One uniform batch through the forward sweep + trace, for an ncu capture of vsk_dp_dir_kernel:
    python scripts/dd_ncu.py [pattern size] [pairs]        (VSK_DD_FORCE selects the plan)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import vntrseek_b200 as v
from vntrseek_b200 import synth

L = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2368
ctx = v.Context(np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "dist_table.npz"))["sm"], 0)
rng = np.random.default_rng(0)
pure, mix = synth._PURE, synth._MIX[synth._MIX > 0]
tot = n * L
ids = pure[rng.integers(0, 4, tot)]
m = rng.random(tot) < 0.15
ids[m] = mix[rng.integers(0, len(mix), int(m.sum()))]
dash = rng.random(tot) < 0.03
ids[dash] = synth._DASH[rng.integers(0, 4, int(dash.sum())), rng.integers(1, 5, int(dash.sum()))]
off = np.arange(n + 1, dtype=np.int64) * L
i2 = ids.copy()
mut = rng.random(tot) < 0.03
i2[mut] = pure[rng.integers(0, 4, int(mut.sum()))]
pa, pb = ctx.profiles(off=off, ids=ids.astype(np.uint16)), ctx.profiles(off=off, ids=i2.astype(np.uint16))
idx = np.arange(n, dtype=np.int32)
b = ctx.dp_batch(pa, pb, idx, idx, np.zeros(n, np.uint8), 88)
for _ in range(2):
    b.run()
ctx.sync()
ctx.timer_start()
b.run()
dt = ctx.timer_stop() * 1e-3
print("L=%d pairs=%d  %.2f ms  %.0f GCUPS" % (L, n, dt * 1e3, b.cells / dt / 1e9))
