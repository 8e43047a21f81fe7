"""Best-reference table from flankalign's .map files.  Host code, no GPU.
The checker is oracle/rankflank_ref.pl: the reference's own scoring loop
(perl/run_rankflankmap.pl:96-231) restated in Perl with the database inserts turned into prints,
so split / regex / double arithmetic / number printing are Perl's; its output for the fixture is
committed (tests/golden/rankflank_small), and it is run live where perl is installed."""
import filecmp
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

PERL = shutil.which("perl")
SCRIPT = os.path.join(ROOT, "oracle", "rankflank_ref.pl")


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


def test_rankflank_bad_args(tmp_path):
    import vntrseek_b200 as v
    with pytest.raises(v.VskError):
        v.rankflank(str(tmp_path / "nope"), str(tmp_path / "r.tsv"))
