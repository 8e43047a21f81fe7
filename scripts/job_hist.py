import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import vntrseek_b200 as v
from vntrseek_b200 import synth
nref, nreads = int(sys.argv[1]), int(sys.argv[2])
os.environ["VSK_DUMP_JOBS"] = "/tmp/jobs.bin"
if os.path.exists("/tmp/jobs.bin"): os.unlink("/tmp/jobs.bin")
table = np.load("tests/golden/dist_table.npz")["sm"]
ctx = v.Context(table, 0)
refs = synth.fast_refs(nref, seed=1); reads = synth.fast_reads(refs, nreads, seed=2)
rt = v.TRSetHandle.from_packed(refs, False); qt = v.TRSetHandle.from_packed(reads, True)
idx = v.RefIndex(ctx, rt)
links, st = v.proclu(ctx, idx, qt)
print(st)
j = np.fromfile("/tmp/jobs.bin", dtype=np.int32).reshape(-1, 3)
l1, l2, br = j[:, 0].astype(np.int64), j[:, 1].astype(np.int64), j[:, 2]
cells = l1 * l2
print("pairs", len(j), "cells", cells.sum(), "banded frac", (br >= 0).mean())
edges = [0, 32, 64, 128, 192, 256, 384, 512, 768, 1024, 1536, 2048, 8192]
for a, b in zip(edges[:-1], edges[1:]):
    m = (l2 > a) & (l2 <= b)
    print(f"l2 in ({a},{b}]: pairs {m.mean()*100:6.2f}%  cells {cells[m].sum()/cells.sum()*100:6.2f}%  mean l1 {l1[m].mean() if m.any() else 0:.0f}")
os.makedirs("gpurun_out", exist_ok=True)
np.save("gpurun_out/jobs_sample.npy", j[:: max(1, len(j) // 400000)])
print("TOTALJOBS", len(j))
