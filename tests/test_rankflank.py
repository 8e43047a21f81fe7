"""Best-reference table from flankalign's .map files.  Host code, no GPU.
Parity is pinned to the reference's own code: oracle/rankflank_harness.pl evaluates the UNMODIFIED
scoring loop of perl/run_rankflankmap.pl:97-231 (its text read from /root/reference, the database
helpers stubbed); the committed fixture tests/golden/rankflank_small was recorded from it, and it is
run live where the reference tree and perl exist.  oracle/rankflank_ref.pl is a restatement of the
same loop for machines without the reference tree (e.g. the GPU box); both are compared here."""
import filecmp
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

PERL = shutil.which("perl")
SCRIPT = os.path.join(ROOT, "oracle", "rankflank_ref.pl")
HARNESS = os.path.join(ROOT, "oracle", "rankflank_harness.pl")
REF_SCRIPT = "/root/reference/perl/run_rankflankmap.pl"


def make_maps(path, seed, nfiles=3, nlines=60):
    """.map files as flankalign writes them (read, dir, leftlen, rightlen, 3 counters, hits) with
    ties, several hits per read, hits under and exactly at 0.90, reads without flanks, hits with
    a tail after the third number, fields Perl's regex rejects, short lines, a `_map*` file"""
    rng = np.random.default_rng(seed)
    os.makedirs(path, exist_ok=True)
    rid = 100
    for f in range(nfiles):
        lines = ["@%d_1\t1\t1\t1\t1\t1" % f]
        for _ in range(nlines):
            rid += 1
            ll, rl = int(rng.integers(0, 51)), int(rng.integers(0, 51))
            if rng.random() < 0.05:
                ll = rl = 0
            hits = []
            for _ in range(int(rng.integers(0, 5))):
                ref = int(rng.integers(1, 40))
                e1, e2 = int(rng.integers(0, 4)), int(rng.integers(0, 4))
                h = "%s%d:%d:%d" % ("-" if rng.random() < 0.8 else "", ref, e1, e2)
                if rng.random() < 0.1:
                    h += ":7x"
                if rng.random() < 0.05:
                    h = "x" + h
                hits.append(h)
            if hits and rng.random() < 0.3:
                hits.append(hits[0])  # an exact tie
            lines.append("%d\t'\t%d\t%d\t1\t1\t1\t%s" % (rid, ll, rl, ",".join(hits)))
        lines.append("%d\t'\t10\t10\t1\t1\t1\t-5:1:1" % (rid + 1000))  # (20-2)/20 = 0.9 exactly
        lines.append("%d\t'\t50\t50\t1\t1\t1\t-5:5:5,-6:6:4,-7:11:0" % (rid + 2000))  # 0.9, 0.9 kept; 0.89 not
        with open(os.path.join(path, "%d_1.map" % f), "w") as fh:
            fh.write("\n".join(lines) + "\n")
    with open(os.path.join(path, "_map_skipped"), "w") as fh:
        fh.write("1\t'\t10\t10\t1\t1\t1\t-5:0:0\n")


def _perl(mapdir, out):
    subprocess.run([PERL, SCRIPT, mapdir, os.path.join(out, "ref_rank.tsv"), os.path.join(out, "ref_map.tsv")],
                   check=True)


def test_rankflank_golden(tmp_path):
    import vntrseek_b200 as v
    g = os.path.join(GOLDEN, "rankflank_small")
    st = v.rankflank(os.path.join(g, "maps"), str(tmp_path / "rank.tsv"), str(tmp_path / "map.tsv"))
    assert filecmp.cmp(str(tmp_path / "rank.tsv"), os.path.join(g, "rank.tsv"), shallow=False)
    assert filecmp.cmp(str(tmp_path / "map.tsv"), os.path.join(g, "map.tsv"), shallow=False)
    assert st["files"] == 3 and st["rank_rows"] == sum(1 for _ in open(os.path.join(g, "rank.tsv")))
    assert st["map_rows"] == sum(1 for _ in open(os.path.join(g, "map.tsv")))
    # the 0.90 boundary and the tie rules, spelled out
    rows = [l.rstrip("\n").split("\t") for l in open(str(tmp_path / "rank.tsv"))]
    exact = [r for r in rows if r[2] == "0.9"]
    assert any(r[0] == "5" and r[3] == "0" for r in exact)                # 1 - 2/20
    assert {r[0] for r in exact if r[3] == "1"} >= {"5", "6"}            # two hits at 0.9, 0.89 dropped
    assert all(float(r[2]) >= 0.9 for r in rows)
    # without the map table
    v.rankflank(os.path.join(g, "maps"), str(tmp_path / "rank2.tsv"))
    assert filecmp.cmp(str(tmp_path / "rank2.tsv"), os.path.join(g, "rank.tsv"), shallow=False)


@pytest.mark.skipif(PERL is None, reason="perl not installed")
@pytest.mark.parametrize("seed,nfiles,nlines", [(1, 1, 10), (2, 4, 300), (3, 7, 2000)])
def test_rankflank_matches_perl(tmp_path, seed, nfiles, nlines):
    import vntrseek_b200 as v
    w = str(tmp_path)
    make_maps(os.path.join(w, "maps"), seed, nfiles, nlines)
    _perl(os.path.join(w, "maps"), w)
    v.rankflank(os.path.join(w, "maps"), os.path.join(w, "rank.tsv"), os.path.join(w, "map.tsv"))
    assert filecmp.cmp(os.path.join(w, "rank.tsv"), os.path.join(w, "ref_rank.tsv"), shallow=False)
    assert filecmp.cmp(os.path.join(w, "map.tsv"), os.path.join(w, "ref_map.tsv"), shallow=False)


@pytest.mark.skipif(PERL is None, reason="perl not installed")
def test_rankflank_on_flankalign_fixture(tmp_path):
    """the .map files of the flankalign fixture (written by the reference flankalign.exe)"""
    import vntrseek_b200 as v
    w = str(tmp_path)
    maps = os.path.join(GOLDEN, "flankalign_small", "expected")
    _perl(maps, w)
    st = v.rankflank(maps, os.path.join(w, "rank.tsv"), os.path.join(w, "map.tsv"))
    assert st["rank_rows"] > 50
    assert filecmp.cmp(os.path.join(w, "rank.tsv"), os.path.join(w, "ref_rank.tsv"), shallow=False)
    assert filecmp.cmp(os.path.join(w, "map.tsv"), os.path.join(w, "ref_map.tsv"), shallow=False)


def make_edge_maps(path):
    """edge cases written out by hand; the expected rows below are derived from the reference's loop
    (run_rankflankmap.pl:109-166) line by line"""
    os.makedirs(path, exist_ok=True)
    lines = [
        "@7_1\t1\t1\t1\t1\t1",                        # header: 6 fields -> skipped (:111)
        "1\t'\t10\t10\t1\t1\t1\t-5:1:1",              # 1-2/20 = 0.9 exactly: kept (>= 0.90)
        "2\t'\t10\t11\t1\t1\t1\t-5:1:1,-6:2:0",       # 1-2/21 twice: tie, both refs, ties=1
        "3\t'\t0\t0\t1\t1\t1\t-5:0:0",                # no flanks: score 0 -> below 0.90 -> map row only
        "4\t'\t30\t20\t1\t1\t1\t-9:3:3,-8:0:0,-7:0:0,-8:0:0",  # 0.88 dropped; 1.0 three times (8,7,8)
        "5\t'\t50\t50\t1\t1\t1\t-5:5:6",              # 0.89: dropped, no rank row
        "6\t'\t3\t0\t1\t1\t1\t12:0:0:junk,x13:0:0,14:0",  # unsigned id, tail ignored; two fields rejected by the regex
        "7\t'\t7\t0\t1\t1\t1\t",                      # no hits: split drops the empty field -> 7 fields -> skipped
        "8\t'\t3\t0\t1\t1\t1\t-21:0:0",               # 1 - 0/3 = 1 (prints as 1)
        "9\t'\t17\t13\t1\t1\t1\t-22:1:1",             # 1 - 2/30 = 0.933333333333333 (%.15g)
    ]
    with open(os.path.join(path, "7_1.map"), "w") as fh:
        fh.write("\n".join(lines) + "\n")


EDGE_RANK = [("5", "1", "0.9", "0"), ("5", "2", "0.904761904761905", "1"), ("6", "2", "0.904761904761905", "1"),
             ("8", "4", "1", "2"), ("7", "4", "1", "2"), ("8", "4", "1", "2"), ("12", "6", "1", "0"),
             ("21", "8", "1", "0"), ("22", "9", "0.933333333333333", "0")]
EDGE_MAP = [("5", "1"), ("5", "2"), ("6", "2"), ("5", "3"), ("9", "4"), ("8", "4"), ("7", "4"), ("8", "4"), ("5", "5"),
            ("12", "6"), ("21", "8"), ("22", "9")]


def test_rankflank_hand_derived_edge_cases(tmp_path):
    """ties, zero-length flanks, a score of exactly 0.90, %.15g printing, fields the regex rejects"""
    import vntrseek_b200 as v
    w = str(tmp_path)
    make_edge_maps(os.path.join(w, "maps"))
    st = v.rankflank(os.path.join(w, "maps"), os.path.join(w, "rank.tsv"), os.path.join(w, "map.tsv"))
    rank = [tuple(l.rstrip("\n").split("\t")) for l in open(os.path.join(w, "rank.tsv"))]
    maps = [tuple(l.rstrip("\n").split("\t")) for l in open(os.path.join(w, "map.tsv"))]
    assert rank == EDGE_RANK
    assert maps == EDGE_MAP
    assert st["rank_rows"] == len(EDGE_RANK) and st["map_rows"] == len(EDGE_MAP)
    if PERL and os.path.exists(REF_SCRIPT):   # and the reference's own loop says the same
        subprocess.run([PERL, HARNESS, REF_SCRIPT, os.path.join(w, "maps"), os.path.join(w, "h_rank.tsv"),
                        os.path.join(w, "h_map.tsv")], check=True)
        assert filecmp.cmp(os.path.join(w, "rank.tsv"), os.path.join(w, "h_rank.tsv"), shallow=False)
        assert filecmp.cmp(os.path.join(w, "map.tsv"), os.path.join(w, "h_map.tsv"), shallow=False)


@pytest.mark.skipif(PERL is None or not os.path.exists(REF_SCRIPT), reason="needs perl and /root/reference")
@pytest.mark.parametrize("seed,nfiles,nlines", [(11, 2, 50), (12, 5, 1500)])
def test_rankflank_matches_unmodified_reference_loop(tmp_path, seed, nfiles, nlines):
    """the library and the Perl restatement against the reference's own loop text, evaluated by
    oracle/rankflank_harness.pl"""
    import vntrseek_b200 as v
    w = str(tmp_path)
    make_maps(os.path.join(w, "maps"), seed, nfiles, nlines)
    subprocess.run([PERL, HARNESS, REF_SCRIPT, os.path.join(w, "maps"), os.path.join(w, "h_rank.tsv"),
                    os.path.join(w, "h_map.tsv")], check=True)
    _perl(os.path.join(w, "maps"), w)
    v.rankflank(os.path.join(w, "maps"), os.path.join(w, "rank.tsv"), os.path.join(w, "map.tsv"))
    for a in ("rank", "map"):
        assert filecmp.cmp(os.path.join(w, a + ".tsv"), os.path.join(w, "h_" + a + ".tsv"), shallow=False)
        assert filecmp.cmp(os.path.join(w, "ref_" + a + ".tsv"), os.path.join(w, "h_" + a + ".tsv"), shallow=False)


def test_rankflank_bad_args(tmp_path):
    import vntrseek_b200 as v
    with pytest.raises(v.VskError):
        v.rankflank(str(tmp_path / "nope"), str(tmp_path / "r.tsv"))
