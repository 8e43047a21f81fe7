"""join_clusters: the merge of per-file PROCLU cluster files.  Host code, so
these tests need no GPU: the committed fixture (inputs + the reference's output), live runs of
the reference binary (oracle/_ref/join_clusters.exe, compiled from the unmodified sources) on
seeded folders, the reference's error exits, and the properties of the merge."""
import filecmp
import importlib.util
import os
import re
import shutil

import pytest

from _cli import OUR_BIN, REF_BIN, run
from conftest import GOLDEN

SUMMARY = re.compile(r"Total clusters read: (\d+), after join: (\d+), total refs\(written\) (\d+) \((\d+)\),  "
                     r"total reads\(written\) (\d+) \((\d+)\) maxdepth=(\d+)")


def _gen():
    spec = importlib.util.spec_from_file_location("make_join_golden", os.path.join(GOLDEN, "make_join_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _have_ref():
    return os.path.exists(os.path.join(REF_BIN, "join_clusters.exe"))


def _summary(text):
    m = SUMMARY.search(text)
    return tuple(int(x) for x in m.groups()) if m else None


def test_join_golden(tmp_path):
    """committed fixture: drop-in binary and library call give the reference's bytes and counters"""
    import vntrseek_b200 as v
    g = os.path.join(GOLDEN, "join_small")
    w = str(tmp_path / "w")
    shutil.copytree(os.path.join(g, "in"), os.path.join(w, "in"))
    rc, out, _ = run(os.path.join(OUR_BIN, "join_clusters.exe"), ["in", "all.clusters"], w)
    assert rc == 0
    assert filecmp.cmp(os.path.join(w, "all.clusters"), os.path.join(g, "expected.out"), shallow=False)
    exp = _summary(open(os.path.join(g, "ref.log")).read())
    assert _summary(out) == exp
    st = v.join_clusters(os.path.join(w, "in"), os.path.join(w, "lib.out"))
    assert filecmp.cmp(os.path.join(w, "lib.out"), os.path.join(g, "expected.out"), shallow=False)
    assert (st["clusters_read"], st["clusters_after_join"], st["refs_read"], st["refs_written"], st["reads_read"],
            st["reads_written"], st["maxdepth"]) == exp
    # explicit list in any order: sorted like the reference sorts its file list; x.clu is not in it
    files = sorted(os.listdir(os.path.join(w, "in")), reverse=True)
    files = [os.path.join(w, "in", f) for f in files if len(f) > 6]
    v.join_clusters(files, os.path.join(w, "lib2.out"))
    assert filecmp.cmp(os.path.join(w, "lib2.out"), os.path.join(g, "expected.out"), shallow=False)


@pytest.mark.parametrize("seed,nfiles,nref,lo,hi", [(1, 3, 40, 5, 12), (2, 6, 300, 40, 80), (3, 5, 5000, 100, 200),
                                                    (4, 1, 30, 20, 40), (5, 12, 2000, 150, 300)])
def test_join_matches_reference_live(tmp_path, seed, nfiles, nref, lo, hi):
    """dense reference space (long merge chains, many flips) to sparse (hardly any merge)"""
    if not _have_ref():
        pytest.skip("oracle/_ref/join_clusters.exe not built")
    w = str(tmp_path)
    _gen().make_folder(os.path.join(w, "in"), seed, nfiles, nref, lo, hi)
    rc1, out1, _ = run(os.path.join(REF_BIN, "join_clusters.exe"), ["in", "ref.out"], w)
    rc2, out2, _ = run(os.path.join(OUR_BIN, "join_clusters.exe"), ["in/", "our.out"], w)
    assert rc1 == 0 and rc2 == 0
    assert filecmp.cmp(os.path.join(w, "ref.out"), os.path.join(w, "our.out"), shallow=False)
    assert _summary(out2) == _summary(out1)
    assert out2.count("flipped c:") == out1.count("flipped c:")


def test_join_empty_and_errors(tmp_path):
    """no cluster file: empty output, exit 0, no summary; malformed files: the reference's message, exit 1"""
    w = str(tmp_path)
    os.makedirs(os.path.join(w, "in"))
    rc, out, _ = run(os.path.join(OUR_BIN, "join_clusters.exe"), ["in", "o.out"], w)
    assert rc == 0 and os.path.getsize(os.path.join(w, "o.out")) == 0 and _summary(out) is None
    rc, out, _ = run(os.path.join(OUR_BIN, "join_clusters.exe"), ["in"], w)
    assert rc == 1 and "Usage" in out
    cases = {"@ -1' 5x": "Unexpected value qualifier detected in cluster file!",
             "-1' 5'": "Invalid format detected in cluster file 1!",
             "@ -1' 5' @ x": "Invalid format detected in cluster file 2!",
             "@ -1' 5' , 7'": "Invalid format detected in cluster file 1!"}
    for k, (text, msg) in enumerate(cases.items()):
        d = os.path.join(w, "bad%d" % k)
        os.makedirs(d)
        with open(os.path.join(d, "reads.leb36.clu"), "w") as fh:
            fh.write(text + "\n")
        rc, out, _ = run(os.path.join(OUR_BIN, "join_clusters.exe"), [d, "o.out"], w)
        assert rc == 1 and msg in out, (text, out)
        if _have_ref():
            rc1, out1, _ = run(os.path.join(REF_BIN, "join_clusters.exe"), [d, "r.out"], w)
            assert rc1 == 1 and msg in out1, (text, out1)


def test_join_properties(tmp_path):
    """every read is written exactly once; a reference ends up in exactly one output cluster; joining the
    joined file again changes nothing but the order rules (same clusters)"""
    import vntrseek_b200 as v
    w = str(tmp_path)
    _gen().make_folder(os.path.join(w, "in"), 11, 8, 400, 60, 120)
    st = v.join_clusters(os.path.join(w, "in"), os.path.join(w, "a.out"))
    assert st["reads_written"] == st["reads_read"]
    seen_ref, seen_read, clusters = {}, set(), []
    for ln, line in enumerate(open(os.path.join(w, "a.out"))):
        ids = [int(t[:-1]) for t in line.strip().split(",") if t]
        for i in ids:
            if i < 0:
                assert seen_ref.setdefault(i, ln) == ln
            else:
                assert i not in seen_read
                seen_read.add(i)
        clusters.append(frozenset(ids))
    assert len(clusters) == st["clusters_after_join"]
    # second join over the first one's output (as a PROCLU-style file: '@' + blank separated members)
    os.makedirs(os.path.join(w, "again"))
    with open(os.path.join(w, "again", "joined.leb36.clu"), "w") as fh:
        for line in open(os.path.join(w, "a.out")):
            fh.write("@ " + " ".join(t for t in line.strip().split(",") if t) + "\n")
    st2 = v.join_clusters(os.path.join(w, "again"), os.path.join(w, "b.out"))
    assert st2["clusters_after_join"] == st["clusters_after_join"]
    again = {frozenset(int(t[:-1]) for t in line.strip().split(",") if t) for line in open(os.path.join(w, "b.out"))}
    assert again == set(clusters)
