"""ad-hoc DP kernel throughput probe (not the bench): hg38-shaped pair lengths."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import vntrseek_b200 as v

table = np.load("tests/golden/dist_table.npz")["sm"]
ctx = v.Context(table, 0)
rng = np.random.default_rng(0)
npairs = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
fixed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
u = rng.random(npairs)
L = np.where(u < .55, rng.integers(7, 31, npairs), np.where(u < .85, rng.integers(31, 101, npairs),
    np.where(u < .97, rng.integers(101, 401, npairs), rng.integers(401, 2001, npairs))))
if fixed: L[:] = fixed
pure = np.array([v.synth.comp_id(10,0,0,0), v.synth.comp_id(0,10,0,0), v.synth.comp_id(0,0,10,0), v.synth.comp_id(0,0,0,10)]) if hasattr(v,'synth') else None
from vntrseek_b200 import synth
pure = np.array([synth.comp_id(10,0,0,0), synth.comp_id(0,10,0,0), synth.comp_id(0,0,10,0), synth.comp_id(0,0,0,10)])
mix = np.array([synth.comp_id(*c) for c in [(9,1,0,0),(8,0,2,0),(7,0,0,3),(0,9,1,0),(1,0,9,0),(0,2,0,8),(5,5,0,0),(0,0,7,0),(0,8,0,0),(3,0,0,7)]])
profs = []
for l in L:
    a = pure[rng.integers(0, 4, l)]
    m = rng.random(l) < 0.15
    a[m] = mix[rng.integers(0, len(mix), int(m.sum()))]
    b = a.copy()
    m = rng.random(l) < 0.03
    b[m] = pure[rng.integers(0, 4, int(m.sum()))]
    profs += [a, b]
ps = ctx.profiles(profs)
li = np.arange(0, 2 * npairs, 2); ti = li + 1
for name, mode in (("full", np.zeros(npairs, np.uint8)), ("narrowband", (L >= 30).astype(np.uint8))):
    b = ctx.dp_batch(ps, ps, li, ti, mode, 88)
    for _ in range(2): b.run()
    ctx.sync()
    t = time.time(); reps = 3
    for _ in range(reps): b.run()
    ctx.sync(); dt = (time.time() - t) / reps
    print(f"{name}: pairs={npairs} cells={b.cells/1e9:.3f}G time={dt*1e3:.2f} ms  GCUPS={b.cells/dt/1e9:.1f}  pairs/s={npairs/dt:.0f}")
    b.close()
