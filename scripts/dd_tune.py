"""Tuning aid: GCUPS of the forward sweep + trace (vsk_dp_batch) per pattern size for every forced plan
(VSK_DD_FORCE = qp16 / qp12 / qp8 / t16 / t12 / t8; unset = the library's own choice).  One process per
setting (the switch is read once).  python scripts/dd_tune.py [sizes...]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, numpy as np
sys.path.insert(0, %r)
import vntrseek_b200 as v
from vntrseek_b200 import synth
sizes = [int(x) for x in sys.argv[1:]]
ctx = v.Context(np.load(%r)["sm"], 0)
rng = np.random.default_rng(0)
pure, mix = synth._PURE, synth._MIX[synth._MIX > 0]
for L in sizes:
    n = int(max(2368, min(100000, 3e9 / (L * L))))
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
    out = []
    for mode in (0, 1):
        b = ctx.dp_batch(pa, pb, idx, idx, np.full(n, mode, np.uint8), 88)
        b.run(); ctx.sync()
        ctx.timer_start()
        for _ in range(3): b.run()
        dt = ctx.timer_stop() / 3 * 1e-3
        out.append(b.cells / dt / 1e9)
        b.close()
    pa.close(); pb.close()
    print("%%d:%%.0f/%%.0f" %% (L, out[0], out[1]), end="  ", flush=True)
print()
''' % (ROOT, os.path.join(ROOT, "tests", "golden", "dist_table.npz"))

sizes = sys.argv[1:] or ["134", "201", "303", "474", "715", "1111", "1500", "2000"]
for force in [""] + (os.environ.get("DD_TUNE_FORCES", "16 12 8").split()):
    env = dict(os.environ)
    if force:
        env["VSK_DD_FORCE"] = force
    print("%-5s" % (force or "auto"), end=" ", flush=True)
    subprocess.run([sys.executable, "-c", CHILD] + sizes, env=env)
