"""GPU, file level, at the sizes the headline numbers are quoted on.

* the Watson-shaped parity configuration (66 read TRs x 230,306 reference TRs):
  the drop-in psearch.exe against what the unmodified reference wrote for the same files
  (tests/golden/watson_shape.json, made by tests/golden/make_watson.py), normal mode and `-r 50 -m`;
* a run cut into many chunks and many direction-word batches against the reference binary;
* the reference's abort conditions (duplicate key, pattern size outside 7..5000);
* two psearch.exe processes sharing one GPU (run_proclu.pl starts NPROC of them).
"""
import json
import os
import shutil
import subprocess
import sys

import pytest

from _cli import OUR_BIN, REF_BIN, canonical_clu, parse_counters, run
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def test_watson_shaped_config_matches_reference(tmp_path, dst_file):
    fixture = os.path.join(GOLDEN, "watson_shape.json")
    want = json.load(open(fixture))
    sys.path.insert(0, GOLDEN)
    import make_watson
    w = str(tmp_path)
    sha = make_watson.make_inputs(w)
    assert sha == want["sha256"], "the generator no longer makes the files the fixture was recorded on: re-run tests/golden/make_watson.py"
    exe = os.path.join(OUR_BIN, "psearch.exe")
    rc, _, log = run(exe, ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50], w)
    assert rc == 0, log[-500:]
    assert parse_counters(log) == want["counters"]
    got = [[list(m) for m in c] for c in canonical_clu(os.path.join(w, "reads.leb36.clu"))]
    assert got == want["clu"]
    assert len(got) > 10
    os.remove(os.path.join(w, "reads.leb36.clu"))
    rc, _, log = run(exe, ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50, "-r", 50, "-m"], w)
    assert rc == 0, log[-500:]
    assert parse_counters(log) == want["r_counters"]
    assert sorted(sorted(line.split()[1:]) for line in open(os.path.join(w, "reads.leb36.clu"))) == want["r_clu"]
    assert sorted(open(os.path.join(w, "reads.leb36.map")).read().splitlines()) == want["r_map"]


def _make(tmp, nref, nread, seed, max_patlen):
    from vntrseek_b200 import synth
    refs = synth.make_refs(nref, seed=seed, max_patlen=max_patlen, rot_frac=0.03)
    reads = synth.make_reads(refs, nread, seed=seed + 100, rot_frac=0.02)
    refs.write_leb36(os.path.join(tmp, "reference.leb36"))
    reads.write_leb36(os.path.join(tmp, "reads.leb36"))
    return refs, reads


def test_many_chunks_and_batches_match_reference(tmp_path, dst_file):
    """key budget of 20,000 candidate keys (dozens of chunks) and 16 MiB of direction words per batch
    (several batches per chunk, both side streams): same counters and clusters as the reference binary"""
    if not os.path.exists(os.path.join(REF_BIN, "psearch.exe")):
        pytest.skip("oracle/_ref/psearch.exe not built")
    a, b = str(tmp_path / "ref"), str(tmp_path / "our")
    os.makedirs(a)
    _make(a, 1500, 2500, 17, 2000)
    shutil.copytree(a, b)
    args = ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50]
    rc1, _, log1 = run(os.path.join(REF_BIN, "psearch.exe"), args, a)
    env = dict(os.environ, VSK_KEY_BUDGET="20000", VSK_DIR_BATCH_MB="16")
    rc2, _, log2 = run(os.path.join(OUR_BIN, "psearch.exe"), args, b, env=env)
    assert rc1 == 0 and rc2 == 0, (rc1, rc2, log2[-500:])
    c = parse_counters(log2)
    assert c == parse_counters(log1)
    assert c["ProfileAlignments"] > 1000
    assert canonical_clu(os.path.join(b, "reads.leb36.clu")) == canonical_clu(os.path.join(a, "reads.leb36.clu"))


def test_duplicate_key_aborts_like_reference(tmp_path, dst_file):
    """psearch.c:397-401,424-428: a key may occur once across both files"""
    w = str(tmp_path)
    refs, reads = _make(w, 60, 80, 3, 100)
    lines = open(os.path.join(w, "reads.leb36")).read().splitlines()
    dup = lines[5].split(" ", 1)
    lines.append(lines[0].split(" ", 1)[0] + " " + dup[1])   # a second record under the first record's key
    open(os.path.join(w, "reads.leb36"), "w").write("\n".join(lines) + "\n")
    args = ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50]
    rc, out, log = run(os.path.join(OUR_BIN, "psearch.exe"), args, w)
    assert rc == 1 and "duplicate index" in (out + log)
    if os.path.exists(os.path.join(REF_BIN, "psearch.exe")):
        rc_ref, out_ref, log_ref = run(os.path.join(REF_BIN, "psearch.exe"), args, w)
        assert rc_ref == 1 and "duplicate index" in (out_ref + log_ref)


@pytest.mark.parametrize("which,patlen", [("reference.leb36", 6), ("reads.leb36", 5001), ("reads.leb36", 3)])
def test_pattern_size_outside_7_5000_aborts_like_reference(tmp_path, dst_file, which, patlen):
    """psearch.c:518-521 (references) and :784-787 (reads): pmin must be in 7..5000"""
    w = str(tmp_path)
    _make(w, 60, 80, 4, 100)
    path = os.path.join(w, which)
    lines = open(path).read().splitlines()
    f = lines[7].split(" ")
    f[1] = str(patlen)
    lines[7] = " ".join(f)
    open(path, "w").write("\n".join(lines) + "\n")
    args = ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50]
    rc, out, log = run(os.path.join(OUR_BIN, "psearch.exe"), args, w)
    assert rc == 1 and "7-5000" in (out + log)
    if os.path.exists(os.path.join(REF_BIN, "psearch.exe")):
        rc_ref, out_ref, log_ref = run(os.path.join(REF_BIN, "psearch.exe"), args, w)
        assert rc_ref == 1 and "7-5000" in (out_ref + log_ref)


def test_two_psearch_processes_share_one_gpu(tmp_path, dst_file):
    """run_proclu.pl keeps NPROC psearch.exe running at once (perl/run_proclu.pl:95-125): two of ours
    on the same device, at the same time, each give the output of a run on its own"""
    from vntrseek_b200 import synth
    refs = synth.make_refs(1200, seed=41, max_patlen=1500, rot_frac=0.0)
    reads = synth.make_reads(refs, 2400, seed=141, rot_frac=0.0)
    half = len(reads.key) // 2
    dirs = []
    for part, (lo, hi) in enumerate(((0, half), (half, len(reads.key)))):
        w = str(tmp_path / ("p%d" % part))
        os.makedirs(w)
        refs.write_leb36(os.path.join(w, "reference.leb36"))
        reads.subset(range(lo, hi)).write_leb36(os.path.join(w, "reads.leb36"))
        dirs.append(w)
    args = [os.path.join(OUR_BIN, "psearch.exe"), "reference.leb36", "reads.leb36", dst_file, "88", "0", "50"]
    solo = []
    for w in dirs:   # one after the other
        p = subprocess.run(args, cwd=w, capture_output=True, text=True)
        assert p.returncode == 0, p.stderr[-500:]
        solo.append((parse_counters(p.stderr), open(os.path.join(w, "reads.leb36.clu")).read()))
        os.remove(os.path.join(w, "reads.leb36.clu"))
    procs = [subprocess.Popen(args, cwd=w, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for w in dirs]
    outs = [p.communicate() for p in procs]
    for p, (so, se), w, want in zip(procs, outs, dirs, solo):
        assert p.returncode == 0, se[-500:]
        assert (parse_counters(se), open(os.path.join(w, "reads.leb36.clu")).read()) == want
    assert solo[0][0]["ProfileAlignments"] > 100
