""".leb36 files of more than a few MB are cut at line starts and parsed by several host threads
(vsk_trset_load_leb36, VSK_LOAD_THREADS).  The threaded load must give exactly what the serial
parser gives -- also where the reference's loader stops at a record it cannot parse
(psearch.c:332-345) and for a file whose records do not sit one per line."""
import os

import pytest


@pytest.fixture(scope="module")
def big_file(tmp_path_factory):
    from vntrseek_b200 import synth
    d = tmp_path_factory.mktemp("leb")
    refs = synth.make_refs(300, seed=3, max_patlen=500)
    reads = synth.make_reads(refs, 12000, seed=4, rot_frac=0.0)
    p = str(d / "reads.leb36")
    reads.write_leb36(p)
    assert os.path.getsize(p) > (4 << 20)
    return p


def _load(path, threads):
    import vntrseek_b200 as v
    from vntrseek_b200._lib import lib
    old = os.environ.get("VSK_LOAD_THREADS")
    os.environ["VSK_LOAD_THREADS"] = str(threads)
    try:
        h = v.TRSetHandle.from_file(path, is_read=True, maxflank=50, rotindex=False)
    finally:
        if old is None:
            del os.environ["VSK_LOAD_THREADS"]
        else:
            os.environ["VSK_LOAD_THREADS"] = old
    return lib().vsk_trset_size(h.h), lib().vsk_trset_checksum(h.h)


def test_threaded_load_equals_serial(big_file):
    want = _load(big_file, 1)
    assert want[0] == 12000
    for t in (2, 3, 4, 8):
        assert _load(big_file, t) == want, t


@pytest.mark.parametrize("where", [0.02, 0.5, 0.97])
def test_stops_at_the_first_bad_record_like_the_serial_parser(big_file, tmp_path, where):
    lines = open(big_file).read().split("\n")
    k = int(len(lines) * where)
    toks = lines[k].split(" ")
    toks[3] = "x"                      # profile length that does not parse: the loader stops here
    lines[k] = " ".join(toks)
    p = str(tmp_path / "bad.leb36")
    open(p, "w").write("\n".join(lines))
    want = _load(p, 1)
    assert want[0] == k
    for t in (2, 4, 8):
        assert _load(p, t) == want, t


def test_ragged_file_falls_back_to_serial(big_file, tmp_path):
    """records broken over two lines, two records on one line: line starts are not record starts"""
    lines = open(big_file).read().split("\n")
    out = []
    for i, line in enumerate(lines):
        if i % 7 == 3 and line:
            toks = line.split(" ")
            out.append(" ".join(toks[:6]))
            out.append(" ".join(toks[6:]))
        elif i % 7 == 5 and line and out:
            out[-1] = out[-1] + "  " + line
        else:
            out.append(line)
    p = str(tmp_path / "ragged.leb36")
    open(p, "w").write("\n".join(out))
    want = _load(p, 1)
    assert want == _load(big_file, 1)   # the same records, only laid out differently
    for t in (2, 4, 8):
        assert _load(p, t) == want, t


def test_bad_composition_id_is_an_error_either_way(big_file, tmp_path):
    import vntrseek_b200 as v
    lines = open(big_file).read().split("\n")
    k = len(lines) // 2
    toks = lines[k].split(" ")
    toks[5] = "ZZ" + toks[5][2:]       # 35 + 36*35 = 1295 > 1000
    lines[k] = " ".join(toks)
    p = str(tmp_path / "badid.leb36")
    open(p, "w").write("\n".join(lines))
    for t in (1, 4):
        with pytest.raises(v.VskError):
            _load(p, t)


def test_packed_cache_serves_the_same_set(tmp_path):
    """vsk_trset_load_leb36_cached: the first load parses the text and writes the binary file, the second
    reads it back (same checksum, also after the rotindex), a changed text file or other parse parameters
    are not served from it, a damaged file is ignored and replaced"""
    import os
    import time

    import vntrseek_b200 as v
    from vntrseek_b200 import synth
    refs = synth.make_refs(300, seed=8, max_patlen=400, rot_frac=0.05)
    reads = synth.make_reads(refs, 500, seed=108, rot_frac=0.03)
    path = str(tmp_path / "reads.leb36")
    reads.write_leb36(path)
    cache = path + ".vsktr"
    plain = v.TRSetHandle.from_file(path, is_read=True, maxflank=50)
    a = v.TRSetHandle.from_file(path, is_read=True, maxflank=50, cache_path=cache)
    assert a.from_cache is False and os.path.getsize(cache) > 1000
    b = v.TRSetHandle.from_file(path, is_read=True, maxflank=50, cache_path=cache)
    assert b.from_cache is True
    assert plain.checksum() == a.checksum() == b.checksum()
    assert len(b) == 500
    # other parse parameters: the cut of the read flanks is baked into the arrays
    c = v.TRSetHandle.from_file(path, is_read=True, maxflank=20, cache_path=cache)
    assert c.from_cache is False
    assert c.checksum() == v.TRSetHandle.from_file(path, is_read=True, maxflank=20).checksum() != plain.checksum()
    # the text changes (one record less): not served
    time.sleep(0.01)
    lines = open(path).read().splitlines()
    open(path, "w").write("\n".join(lines[:-1]) + "\n")
    d = v.TRSetHandle.from_file(path, is_read=True, maxflank=20, cache_path=cache, rotindex=False)
    assert d.from_cache is False and len(d) == 499
    # damaged file
    blob = open(cache, "rb").read()
    open(cache, "wb").write(blob[: len(blob) // 2])
    e = v.TRSetHandle.from_file(path, is_read=True, maxflank=20, cache_path=cache, rotindex=False)
    assert e.from_cache is False and e.checksum() == d.checksum()
    f = v.TRSetHandle.from_file(path, is_read=True, maxflank=20, cache_path=cache, rotindex=False)
    assert f.from_cache is True and f.checksum() == d.checksum()
    assert not [x for x in os.listdir(str(tmp_path)) if ".tmp" in x]
