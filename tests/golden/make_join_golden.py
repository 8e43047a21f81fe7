"""Makes tests/golden/join_small: a folder of synthetic PROCLU cluster files and what the
UNMODIFIED reference join_clusters (oracle/_ref/join_clusters.exe, built by `make -C oracle ref`)
writes for it.  Run in the build container (needs /root/reference); the fixture is committed."""
import os
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def make_folder(path, seed, nfiles=4, nref=120, lo=15, hi=30):
    """cluster files with shared references (merges and merge chains), direction conflicts (flips),
    a reference repeated inside a cluster, comment tokens and ragged white space"""
    rng = np.random.default_rng(seed)
    os.makedirs(path, exist_ok=True)
    rid = 0
    for f in range(nfiles):
        lines = []
        for _ in range(int(rng.integers(lo, hi))):
            refs = rng.choice(nref, int(rng.integers(1, 4)), replace=False) + 1
            mem = ["-%d%s" % (r, "'" if rng.random() < 0.7 else '"') for r in refs]
            if rng.random() < 0.1:
                mem.append(mem[0])
            for _ in range(int(rng.integers(1, 6))):
                rid += 1
                mem.append("%d%s" % (rid, "'" if rng.random() < 0.5 else '"'))
            rng.shuffle(mem)
            line = "@ " + (" " if rng.random() < 0.8 else "  \t").join(mem)
            if rng.random() < 0.1:
                line += " #note"
            lines.append(line)
        with open(os.path.join(path, "reads%02d.leb36.clu" % f), "w") as fh:
            fh.write("\n".join(lines) + "\n")
    with open(os.path.join(path, "x.clu"), "w") as fh:  # name of 5 characters: the reference skips it
        fh.write("@ -1' 999999'\n")


if __name__ == "__main__":
    exe = os.path.join(ROOT, "oracle", "_ref", "join_clusters.exe")
    if not os.path.exists(exe):
        sys.exit("build oracle/_ref first: make -C oracle ref")
    out = os.path.join(HERE, "join_small")
    shutil.rmtree(out, ignore_errors=True)
    make_folder(os.path.join(out, "in"), seed=7)
    p = subprocess.run([exe, "in", "expected.out"], cwd=out, capture_output=True, text=True, check=True)
    with open(os.path.join(out, "ref.log"), "w") as fh:
        fh.write(p.stdout.splitlines()[-1] + "\n")  # the summary line
