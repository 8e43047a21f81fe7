"""Generates tests/golden/watson_shape.json: the reference's answers on the Watson-shaped parity
configuration (the stand-in for BASELINE.json configs[0], whose inputs --
TRF and the t26__ reference set of 230,306 TRs -- are not in the reference repository):
66 read TRs against 230,306 reference TRs, fixed seeds.

Run where /root/reference exists (oracle/_ref built by `make -C oracle ref`); takes a few minutes
(the reference rebuilds its index in each of the two runs).  The fixture holds the SHA-256 of the
two input files, so the test notices if the generator ever produces other inputs.

    python tests/golden/make_watson.py
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

NREF, NREAD, SEED_REF, SEED_READ = 230306, 66, 11, 12


def make_inputs(workdir):
    """writes reference.leb36 / reads.leb36 (+ .rotindex) and eucledian.dst; returns their SHA-256"""
    import numpy as np

    from vntrseek_b200 import synth
    refs = synth.fast_refs(NREF, seed=SEED_REF)
    reads = synth.fast_reads(refs, NREAD, seed=SEED_READ)
    refs.write_leb36(os.path.join(workdir, "reference.leb36"))
    reads.write_leb36(os.path.join(workdir, "reads.leb36"))
    table = np.load(os.path.join(HERE, "dist_table.npz"))["sm"]
    with open(os.path.join(workdir, "eucledian.dst"), "wb") as f:
        f.write(np.array([777, 10, 5], np.int32).tobytes())
        f.write(table.tobytes())
    sha = {}
    for name in ("reference.leb36", "reads.leb36"):
        h = hashlib.sha256()
        with open(os.path.join(workdir, name), "rb") as f:
            for blk in iter(lambda: f.read(1 << 20), b""):
                h.update(blk)
        sha[name] = h.hexdigest()
    return sha


def main():
    from vntrseek_b200.clu import canonical_clu, parse_counters
    exe = os.path.join(ROOT, "oracle", "_ref", "psearch.exe")
    work = tempfile.mkdtemp(prefix="vsk_watson_")
    sha = make_inputs(work)
    out = {"nref": NREF, "nread": NREAD, "seed_ref": SEED_REF, "seed_read": SEED_READ, "sha256": sha,
           "generator": "tests/golden/make_watson.py", "reference": "oracle/_ref/psearch.exe (unmodified reference)"}
    p = subprocess.run([exe, "reference.leb36", "reads.leb36", "eucledian.dst", "88", "0", "50"], cwd=work,
                       capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-500:]
    out["counters"] = parse_counters(p.stderr)
    out["clu"] = [[list(m) for m in c] for c in canonical_clu(os.path.join(work, "reads.leb36.clu"))]
    os.remove(os.path.join(work, "reads.leb36.clu"))
    p = subprocess.run([exe, "reference.leb36", "reads.leb36", "eucledian.dst", "88", "0", "50", "-r", "50", "-m"],
                       cwd=work, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-500:]
    out["r_counters"] = parse_counters(p.stderr)
    out["r_clu"] = sorted(sorted(line.split()[1:]) for line in open(os.path.join(work, "reads.leb36.clu")))
    out["r_map"] = sorted(open(os.path.join(work, "reads.leb36.map")).read().splitlines())
    with open(os.path.join(HERE, "watson_shape.json"), "w") as f:
        json.dump(out, f, indent=0)
    print(json.dumps({k: out[k] for k in ("counters", "r_counters")}), len(out["clu"]), len(out["r_map"]))


if __name__ == "__main__":
    main()
