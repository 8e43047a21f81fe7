"""helpers for the file-level (drop-in binary) parity tests"""
import os
import subprocess

from conftest import ROOT

REF_BIN = os.path.join(ROOT, "oracle", "_ref")
OUR_BIN = os.path.join(ROOT, "vntrseek_b200", "bin")

from vntrseek_b200.clu import COUNTERS, canonical_clu, parse_counters  # noqa: E402,F401


def run(exe, args, cwd, env=None):
    p = subprocess.run([exe] + [str(a) for a in args], cwd=cwd, capture_output=True, text=True, env=env)
    return p.returncode, p.stdout, p.stderr
