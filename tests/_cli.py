"""helpers for the file-level (drop-in binary) parity tests"""
import os
import re
import subprocess

from conftest import ROOT

REF_BIN = os.path.join(ROOT, "oracle", "_ref")
OUR_BIN = os.path.join(ROOT, "vntrseek_b200", "bin")

COUNTERS = ("CandidatesConsidered", "LCSalignments", "ProfileAlignments", "ProfileAlignmentsSuccess",
            "pairs confirmed after flank LCS alignments", "total clustered")


def parse_counters(log):
    out = {}
    for name in COUNTERS:
        m = re.search(r"\t" + re.escape(name) + r": (\d+)", log)
        out[name] = int(m.group(1)) if m else None
    return out


def canonical_clu(path):
    """clusters -> sorted tuple of ((key, dir), ...) with the smallest key normalised to dir 0"""
    clusters = []
    for line in open(path):
        toks = line.split()
        assert toks and toks[0] == "@", line
        mem = []
        for t in toks[1:]:
            mem.append((int(t[:-1]), {"'": 0, '"': 1, "+": 0, "-": 1}[t[-1]]))
        mem.sort()
        flip = mem[0][1]
        clusters.append(tuple((k, d ^ flip) for k, d in mem))
    clusters.sort()
    return clusters


def run(exe, args, cwd, env=None):
    p = subprocess.run([exe] + [str(a) for a in args], cwd=cwd, capture_output=True, text=True, env=env)
    return p.returncode, p.stdout, p.stderr
