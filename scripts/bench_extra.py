"""Secondary measurements (BASELINE.json configs[3] and [4]): pattern-size sweep of the profile
DP kernels (edges-style pair batches, full matrix and narrowband) and the flank-alignment
pair kernel.  Prints markdown tables; results are copied into profiles/."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import vntrseek_b200 as v
from vntrseek_b200 import synth
from vntrseek_b200._lib import lib

table = np.load("tests/golden/dist_table.npz")["sm"]
ctx = v.Context(table, 0)
rng = np.random.default_rng(0)
pure, mix = synth._PURE, synth._MIX[synth._MIX > 0]


def profiles(lens):
    tot = int(lens.sum())
    ids = pure[rng.integers(0, 4, tot)]
    m = rng.random(tot) < 0.15
    ids[m] = mix[rng.integers(0, len(mix), int(m.sum()))]
    off = np.zeros(len(lens) + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    return off, ids.astype(np.uint16)


print("## Pattern-size sweep (TRANGE boundaries, `psearch.c:286-315`), pairs of equal-size profiles\n")
print("| pattern size | pairs | full: GCUPS | full: pairs/s | narrowband: GCUPS (band cells) | narrowband: pairs/s |")
print("|---|---|---|---|---|---|")
for L in (7, 10, 17, 25, 30, 39, 59, 89, 134, 201, 303, 474, 715, 1111, 2000):
    n = int(max(4000, min(400000, 4e9 / (L * L))))
    lens = np.full(n, L, np.int64)
    o1, i1 = profiles(lens)
    o2, i2 = o1.copy(), i1.copy()
    mut = rng.random(len(i2)) < 0.03
    i2[mut] = pure[rng.integers(0, 4, int(mut.sum()))]
    pa, pb = ctx.profiles(off=o1, ids=i1), ctx.profiles(off=o2, ids=i2)
    idx = np.arange(n, dtype=np.int32)
    row = [str(L), str(n)]
    for mode in (0, 1):
        b = ctx.dp_batch(pa, pb, idx, idx, np.full(n, mode, np.uint8), 88)
        b.run(); ctx.sync()
        ctx.timer_start()
        for _ in range(3):
            b.run()
        dt = ctx.timer_stop() / 3 * 1e-3
        row += ["%.0f" % (b.cells / dt / 1e9), "%.3g" % (n / dt)]
        b.close()
    pa.close(); pb.close()
    print("| " + " | ".join(row) + " |")

print("\n## Flank-alignment pair kernel (`vsk_flankalign_kernel`): reference flanks 60+60 nt, read flanks 10-50 nt\n")
L = lib()
L.vsk_flank_pairs.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32,
                              C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.POINTER(C.c_float)]
nref, nread = 50, 200000
dt_rec = np.dtype([("lo", np.int64), ("ro", np.int64), ("ll", np.int32), ("rl", np.int32)])
acgt = np.frombuffer(b"ACGT", np.uint8)
ref_fl = acgt[rng.integers(0, 4, (nref, 120))]
refs = np.zeros(nref, dt_rec)
refs["lo"] = np.arange(nref) * 120
refs["ro"] = refs["lo"] + 60
refs["ll"] = refs["rl"] = 60
pool = [ref_fl.reshape(-1)]
base = nref * 120
src = rng.integers(0, nref, nread)
ll, rl = rng.integers(10, 51, nread), rng.integers(10, 51, nread)
reads = np.zeros(nread, dt_rec)
reads["lo"] = base + np.arange(nread) * 100
reads["ro"] = reads["lo"] + 50
reads["ll"], reads["rl"] = ll, rl
rd = np.zeros((nread, 100), np.uint8) + ord("A")
rd[:, :50] = ref_fl[src][:, :50]
rd[:, 50:] = ref_fl[src][:, 60:110]
m = rng.random(rd.shape) < 0.03
rd[m] = acgt[rng.integers(0, 4, int(m.sum()))]
pool.append(rd.reshape(-1))
pool = np.concatenate(pool).tobytes()
npairs = 20_000_000
pr = np.repeat(np.arange(nread, dtype=np.int32), npairs // nread)
pf = rng.integers(0, nref, len(pr)).astype(np.int32)
out = np.zeros(len(pr), np.uint32)
ms = C.c_float()
for rep in range(2):
    t = time.perf_counter()
    rc = L.vsk_flank_pairs(ctx.h, pool, len(pool), refs.ctypes.data, nref, reads.ctypes.data, nread, pr.ctypes.data,
                           pf.ctypes.data, len(pr), 0, out.ctypes.data, C.byref(ms))
    wall = time.perf_counter() - t
assert rc == 0
cells = 4.0 * len(pr) * 60 * 30
print("| pairs | kernel ms | pairs/s (kernel) | pairs/s (host buffers in, results out) | bit-GCUPS (kernel) | accepted |")
print("|---|---|---|---|---|---|")
print("| %d | %.2f | %.3g | %.3g | %.0f | %.1f%% |" % (len(pr), ms.value, len(pr) / (ms.value * 1e-3), len(pr) / wall,
                                                    cells / (ms.value * 1e-3) / 1e9, 100.0 * ((out >> 16) & 1).mean()))
