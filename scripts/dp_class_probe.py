"""time the DP kernels per size class on a sample of the pipeline's real job sizes
(run scripts/job_hist.py first: it leaves gpurun_out/jobs_sample.npy)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import vntrseek_b200 as v
from vntrseek_b200 import synth
j = np.load("gpurun_out/jobs_sample.npy")
nreads = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
total_jobs = int(sys.argv[2]) if len(sys.argv) > 2 else len(j)
scale = total_jobs / len(j) * (1e6 / nreads)   # sample -> 1M reads
table = np.load("tests/golden/dist_table.npz")["sm"]
ctx = v.Context(table, 0)
rng = np.random.default_rng(0)
pure = synth._PURE; mix = synth._MIX[synth._MIX > 0]
def profs_for(l):
    tot = int(l.sum())
    ids = pure[rng.integers(0, 4, tot)]
    m = rng.random(tot) < 0.15
    ids[m] = mix[rng.integers(0, len(mix), int(m.sum()))]
    off = np.zeros(len(l) + 1, np.int64); np.cumsum(l, out=off[1:])
    return off, ids.astype(np.uint16)
l1, l2, br = j[:, 0], j[:, 1], j[:, 2]
tiny = (br < 0) & (l2 <= 32) & (l1 <= 64)
classes = {
  "tiny(full,l2<=32,l1<=64)": tiny,
  "small other (l2<=32)": ~tiny & (l2 <= 32),
  "l2 33-64": (l2 > 32) & (l2 <= 64),
  "l2 65-128": (l2 > 64) & (l2 <= 128),
  "l2 129-256": (l2 > 128) & (l2 <= 256),
  "l2 257-1024": (l2 > 256) & (l2 <= 1024),
  "l2 >1024": (l2 > 1024),
}
tot = 0
for name, m in classes.items():
    if not m.any(): continue
    a, b, r = l1[m].astype(np.int64), l2[m].astype(np.int64), br[m]
    o1, i1 = profs_for(a); o2, i2 = profs_for(b)
    pa = ctx.profiles(off=o1, ids=i1); pb = ctx.profiles(off=o2, ids=i2)
    n = len(a)
    idx = np.arange(n, dtype=np.int32)
    bt = ctx.dp_batch(pa, pb, idx, idx, (r >= 0).astype(np.uint8), 88)
    bt.run(); ctx.sync()
    t = time.time(); bt.run(); bt.run(); ctx.sync(); dt = (time.time() - t) / 2
    est = dt * scale
    tot += est
    print(f"{name:28s} pairs {n:7d} ({m.mean()*100:5.1f}%) banded {100*(r>=0).mean():5.1f}%  cells {bt.cells/1e9:7.3f}G  time {dt*1e3:8.2f} ms  GCUPS {bt.cells/dt/1e9:7.1f}  -> est per 1M reads {est*1e3:7.1f} ms")
    bt.close(); pa.close(); pb.close()
print("estimated DP total per 1M reads: %.1f ms" % (tot * 1e3))
