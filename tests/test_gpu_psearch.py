"""GPU, file level: the drop-in psearch.exe against the reference psearch.exe
(oracle/_ref, compiled from the unmodified reference sources) on the same
.leb36 / .rotindex files: the six stderr counters, the canonical .clu, and -- in
edges mode -- the byte-exact .edges file."""
import filecmp
import os
import shutil

import numpy as np
import pytest

from _cli import OUR_BIN, REF_BIN, canonical_clu, parse_counters, run

pytestmark = pytest.mark.gpu


def _have_ref():
    return os.path.exists(os.path.join(REF_BIN, "psearch.exe"))


def _make(tmp, nref, nread, seed, max_patlen=600, rot=True):
    from vntrseek_b200 import synth
    refs = synth.make_refs(nref, seed=seed, max_patlen=max_patlen, rot_frac=0.03 if rot else 0.0)
    reads = synth.make_reads(refs, nread, seed=seed + 100, rot_frac=0.02 if rot else 0.0)
    refs.write_leb36(os.path.join(tmp, "reference.leb36"))
    reads.write_leb36(os.path.join(tmp, "reads.leb36"))
    return refs, reads


@pytest.mark.parametrize("nref,nread,seed,maxlen", [(300, 400, 3, 200), (2000, 3000, 1, 600), (1500, 1500, 7, 2000)])
def test_psearch_matches_reference(tmp_path, dst_file, nref, nread, seed, maxlen):
    if not _have_ref():
        pytest.skip("oracle/_ref/psearch.exe not built")
    a, b = str(tmp_path / "ref"), str(tmp_path / "our")
    os.makedirs(a)
    _make(a, nref, nread, seed, maxlen)
    shutil.copytree(a, b)
    args = ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50]
    rc1, _, log1 = run(os.path.join(REF_BIN, "psearch.exe"), args, a)
    rc2, _, log2 = run(os.path.join(OUR_BIN, "psearch.exe"), args, b)
    assert rc1 == 0 and rc2 == 0, (rc1, rc2, log2[-500:])
    assert parse_counters(log2) == parse_counters(log1)
    assert canonical_clu(os.path.join(b, "reads.leb36.clu")) == canonical_clu(os.path.join(a, "reads.leb36.clu"))


def test_psearch_very_long_patterns(tmp_path, dst_file, monkeypatch):
    """patterns of 2100-4600 columns: more than 32 64-bit blocks in the LCS filter (several
    systolic passes), several column strips and direction-word batches in the profile DP"""
    if not _have_ref():
        pytest.skip("oracle/_ref/psearch.exe not built")
    from vntrseek_b200 import synth
    orig = synth._pattern_length
    monkeypatch.setattr(synth, "_pattern_length",
                        lambda rng: int(rng.integers(2100, 4601)) if rng.random() < 0.25 else orig(rng))
    a, b = str(tmp_path / "ref"), str(tmp_path / "our")
    os.makedirs(a)
    _make(a, 120, 160, 21, 5000)
    shutil.copytree(a, b)
    args = ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50]
    rc1, _, log1 = run(os.path.join(REF_BIN, "psearch.exe"), args, a)
    rc2, _, log2 = run(os.path.join(OUR_BIN, "psearch.exe"), args, b)
    assert rc1 == 0 and rc2 == 0, (rc1, rc2, log2[-500:])
    c = parse_counters(log2)
    assert c == parse_counters(log1)
    assert c["ProfileAlignments"] > 0
    assert canonical_clu(os.path.join(b, "reads.leb36.clu")) == canonical_clu(os.path.join(a, "reads.leb36.clu"))


def _canonical_joined(path):
    """all.clusters -> set of clusters, each a frozenset of (id, dir) with the smallest id's mark normalised"""
    out = set()
    for line in open(path):
        mem = [(int(t[:-1]), t[-1]) for t in line.strip().split(",") if t]
        flip = min(mem)[1] != "'"
        out.add(frozenset((i, ("'" if (d == "'") != flip else '"')) for i, d in mem))
    return out


def test_sharded_psearch_then_join(tmp_path, dst_file):
    """the multi-GPU shape of the step at file level: the reads in two files, one psearch run per
    file, join_clusters over the two .clu -- ours end to end against the reference end to end"""
    if not (_have_ref() and os.path.exists(os.path.join(REF_BIN, "join_clusters.exe"))):
        pytest.skip("oracle/_ref binaries not built")
    from vntrseek_b200 import synth
    refs = synth.make_refs(500, seed=31, max_patlen=300, rot_frac=0.0)
    reads = synth.make_reads(refs, 800, seed=131, rot_frac=0.0)
    half = len(reads.key) // 2
    outs = {}
    for name, bindir in (("ref", REF_BIN), ("our", OUR_BIN)):
        w = str(tmp_path / name)
        os.makedirs(os.path.join(w, "clu"))
        refs.write_leb36(os.path.join(w, "reference.leb36"))
        for part, (a, b) in enumerate(((0, half), (half, len(reads.key)))):
            reads.subset(range(a, b)).write_leb36(os.path.join(w, "reads%d.leb36" % part))
            rc, _, log = run(os.path.join(bindir, "psearch.exe"),
                             ["reference.leb36", "reads%d.leb36" % part, dst_file, 88, 0, 50], w)
            assert rc == 0, log[-500:]
            shutil.move(os.path.join(w, "reads%d.leb36.clu" % part), os.path.join(w, "clu"))
        rc, out, _ = run(os.path.join(bindir, "join_clusters.exe"), ["clu", "all.clusters"], w)
        assert rc == 0
        outs[name] = (_canonical_joined(os.path.join(w, "all.clusters")), out.splitlines()[-1])
    assert len(outs["ref"][0]) > 20
    assert outs["our"][0] == outs["ref"][0]
    assert outs["our"][1] == outs["ref"][1]  # the summary line: clusters read / after join / refs / reads


def test_psearch_golden_small(tmp_path, dst_file):
    """committed fixture (inputs + the reference's outputs), no reference binary needed"""
    from conftest import GOLDEN
    g = os.path.join(GOLDEN, "psearch_small")
    w = str(tmp_path / "w")
    shutil.copytree(g, w)
    os.rename(os.path.join(w, "reads.leb36.clu"), os.path.join(w, "expected.clu"))
    rc, _, log = run(os.path.join(OUR_BIN, "psearch.exe"), ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50], w)
    assert rc == 0, log[-500:]
    assert parse_counters(log) == parse_counters(open(os.path.join(w, "ref.log")).read())
    assert canonical_clu(os.path.join(w, "reads.leb36.clu")) == canonical_clu(os.path.join(w, "expected.clu"))


@pytest.mark.parametrize("switch", ["VSK_DEDUP_BITMAP", "VSK_DP_DIR", "VSK_DP_DIRQ8", "VSK_DP_TINYDIR",
                                    "VSK_DP_SMALL_ORDER"])
def test_alternative_paths_give_the_same_files(tmp_path, dst_file, switch):
    """the run-time switches select the older code paths (radix sort + unique instead of the
    per-profile bitmap; one-kernel DP instead of sweep + trace; small-pair kernels beside the first
    sweeps instead of after the join): same counters, same .clu"""
    from conftest import GOLDEN
    g = os.path.join(GOLDEN, "psearch_small")
    outs = []
    for val in (None, "0"):
        w = str(tmp_path / ("w" + str(val)))
        shutil.copytree(g, w)
        os.remove(os.path.join(w, "reads.leb36.clu"))
        env = dict(os.environ)
        if val is not None:
            env[switch] = val
        rc, _, log = run(os.path.join(OUR_BIN, "psearch.exe"), ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50], w,
                         env=env)
        assert rc == 0, log[-500:]
        outs.append((parse_counters(log), open(os.path.join(w, "reads.leb36.clu")).read()))
    assert outs[0] == outs[1]
    assert outs[0][0] == parse_counters(open(os.path.join(g, "ref.log")).read())


def test_index_cache_same_output(tmp_path, dst_file):
    """VSK_INDEX_CACHE=1: the first run writes <REFPROFILES>.vskidx, the second reads it, a damaged
    file is ignored -- counters and .clu are those of a run without the cache every time"""
    from conftest import GOLDEN
    g = os.path.join(GOLDEN, "psearch_small")
    want = None
    for step in ("plain", "write", "read", "damaged"):
        w = str(tmp_path / step)
        shutil.copytree(g, w)
        os.remove(os.path.join(w, "reads.leb36.clu"))
        env = dict(os.environ)
        idx = os.path.join(w, "reference.leb36.vskidx")
        if step != "plain":
            env["VSK_INDEX_CACHE"] = "1"
        if step == "read":
            shutil.copy(str(tmp_path / "write" / "reference.leb36.vskidx"), idx)
            before = os.path.getmtime(idx)
        if step == "damaged":
            open(idx, "wb").write(open(str(tmp_path / "write" / "reference.leb36.vskidx"), "rb").read()[:5000])
        rc, _, log = run(os.path.join(OUR_BIN, "psearch.exe"), ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50], w,
                         env=env)
        assert rc == 0, log[-500:]
        got = (parse_counters(log), open(os.path.join(w, "reads.leb36.clu")).read())
        want = want or got
        assert got == want, step
        assert os.path.exists(idx) == (step != "plain")
        if step == "read":
            assert os.path.getmtime(idx) == before   # used, not rewritten
        if step == "damaged":
            assert os.path.getsize(idx) > 5000       # replaced by a good one


def test_edges_mode_byte_exact(tmp_path, dst_file):
    if not _have_ref():
        pytest.skip("oracle/_ref/psearch.exe not built")
    a, b = str(tmp_path / "ref"), str(tmp_path / "our")
    os.makedirs(a)
    refs, reads = _make(a, 200, 300, 5, 400, rot=False)
    rng = np.random.default_rng(0)
    with open(os.path.join(a, "x.edgein"), "w") as f:
        for _ in range(500):
            f.write("%d,%d\n" % (refs.key[int(rng.integers(0, len(refs)))], reads.key[int(rng.integers(0, len(reads)))]))
    shutil.copytree(a, b)
    args = ["reference.leb36", "reads.leb36", dst_file, 70, 0, 0, "x.edgein"]
    rc1, _, _ = run(os.path.join(REF_BIN, "psearch.exe"), args, a)
    rc2, _, log2 = run(os.path.join(OUR_BIN, "psearch.exe"), args, b)
    assert rc1 == 0 and rc2 == 0, log2[-500:]
    assert filecmp.cmp(os.path.join(a, "reads.leb36.edges"), os.path.join(b, "reads.leb36.edges"), shallow=False)


def test_cli_errors_match_reference(tmp_path, dst_file):
    exe = os.path.join(OUR_BIN, "psearch.exe")
    assert run(exe, [], str(tmp_path))[0] == 1                      # usage
    assert run(exe, ["a", "b", dst_file, 50, 0, 50], str(tmp_path))[0] == 1   # cutoff outside 70..100
    assert run(exe, ["a", "b", dst_file, 88, 0, 50, "-m"], str(tmp_path))[0] == 1  # -m at argv[7] is taken as edges file
    assert run(exe, ["nofile", "nofile", dst_file, 88, 0, 50], str(tmp_path))[0] == 1


def _canonical_clu_r(path):
    """-r mode: members carry + / - (a property of the key, not a relative direction)"""
    out = []
    for line in open(path):
        toks = line.split()
        assert toks[0] == "@"
        out.append(tuple(sorted((int(t[:-1]), t[-1]) for t in toks[1:])))
    return sorted(out)


@pytest.mark.parametrize("reflen", [50, 20])
def test_psearch_ref_to_ref_mode(tmp_path, dst_file, reflen):
    """`-r LEN -m` (produce_indist.pl:156-160): REFLEN flank cut, both flank orders, leftcon/rightcon
    marks, +/- cluster members, and the map file (compared as a sorted multiset of lines)."""
    if not _have_ref():
        pytest.skip("oracle/_ref/psearch.exe not built")
    a, b = str(tmp_path / "ref"), str(tmp_path / "our")
    os.makedirs(a)
    _make(a, 600, 900, 21, 300)
    shutil.copytree(a, b)
    args = ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50, "-r", reflen, "-m"]
    rc1, _, log1 = run(os.path.join(REF_BIN, "psearch.exe"), args, a)
    rc2, _, log2 = run(os.path.join(OUR_BIN, "psearch.exe"), args, b)
    assert rc1 == 0 and rc2 == 0, (rc1, rc2, log2[-500:])
    assert parse_counters(log2) == parse_counters(log1)
    assert _canonical_clu_r(os.path.join(b, "reads.leb36.clu")) == _canonical_clu_r(os.path.join(a, "reads.leb36.clu"))
    m1 = sorted(open(os.path.join(a, "reads.leb36.map")).read().splitlines())
    m2 = sorted(open(os.path.join(b, "reads.leb36.map")).read().splitlines())
    assert len(m1) > 50 and m1 == m2


@pytest.mark.parametrize("switch", ["VSK_GPU_DECODE", "VSK_TRSET_CACHE"])
def test_loader_switches_give_the_same_files(tmp_path, dst_file, switch):
    """VSK_GPU_DECODE=1 (read file tokenised on the host, decoded on the GPU) and VSK_TRSET_CACHE=1 (parsed
    arrays kept in <file>.vsktr; run twice: written, then read back): same counters, same .clu"""
    from conftest import GOLDEN
    g = os.path.join(GOLDEN, "psearch_small")
    want = None
    for run_no, val in enumerate((None, "1", "1")):
        w = str(tmp_path / "w")
        if run_no == 0 or switch != "VSK_TRSET_CACHE":
            shutil.rmtree(w, ignore_errors=True)
            shutil.copytree(g, w)
        if os.path.exists(os.path.join(w, "reads.leb36.clu")):
            os.remove(os.path.join(w, "reads.leb36.clu"))
        env = dict(os.environ)
        if val is not None:
            env[switch] = val
        rc, _, log = run(os.path.join(OUR_BIN, "psearch.exe"), ["reference.leb36", "reads.leb36", dst_file, 88, 0, 50], w,
                         env=env)
        assert rc == 0, log[-500:]
        got = (parse_counters(log), open(os.path.join(w, "reads.leb36.clu")).read())
        want = want or got
        assert got == want, (switch, run_no)
        if switch == "VSK_TRSET_CACHE" and run_no > 0:
            assert os.path.exists(os.path.join(w, "reads.leb36.vsktr")) and os.path.exists(os.path.join(w, "reference.leb36.vsktr"))
    assert want[0] == parse_counters(open(os.path.join(g, "ref.log")).read())
