"""The on-disk cache of the reference seed index (vsk_refs_create_cached and friends): host side
here (no GPU), the drop-in binary with VSK_INDEX_CACHE in tests/test_gpu_psearch.py."""
import os

import pytest


@pytest.fixture(scope="module")
def sets(tmp_path_factory):
    import vntrseek_b200 as v
    from vntrseek_b200 import synth
    d = tmp_path_factory.mktemp("idx")
    a = synth.make_refs(400, seed=4, max_patlen=300, rot_frac=0.05)
    b = synth.make_refs(400, seed=5, max_patlen=300, rot_frac=0.0)
    a.write_leb36(str(d / "a.leb36"))
    b.write_leb36(str(d / "b.leb36"))
    ha = v.TRSetHandle.from_file(str(d / "a.leb36"), is_read=False)                  # with its rotindex
    ha_norot = v.TRSetHandle.from_file(str(d / "a.leb36"), is_read=False, rotindex=False)
    hb = v.TRSetHandle.from_file(str(d / "b.leb36"), is_read=False)
    return d, ha, ha_norot, hb


def test_build_hit_verify(sets):
    import vntrseek_b200 as v
    d, ha, _, _ = sets
    p = str(d / "a.vskidx")
    assert v.index_cache_build(ha, p) is False      # built and written
    assert os.path.getsize(p) > 1000
    assert v.index_cache_build(ha, p) is True       # read back
    assert v.index_cache_verify(ha, p)              # equals a fresh build, array by array
    assert not [f for f in os.listdir(str(d)) if ".tmp" in f]


def test_other_reference_set_is_not_served(sets):
    """a file made for another set -- or for the same records with other master flags (no rotindex)
    -- is ignored and replaced"""
    import vntrseek_b200 as v
    d, ha, ha_norot, hb = sets
    p = str(d / "shared.vskidx")
    v.index_cache_build(ha, p)
    assert not v.index_cache_verify(hb, p)
    assert v.index_cache_build(hb, p) is False
    assert v.index_cache_verify(hb, p) and not v.index_cache_verify(ha, p)
    assert v.index_cache_build(ha_norot, p) is False
    assert v.index_cache_verify(ha_norot, p) and not v.index_cache_verify(ha, p)


@pytest.mark.parametrize("damage", ["truncate", "magic", "garbage_len", "empty"])
def test_damaged_file_is_rebuilt(sets, damage):
    import vntrseek_b200 as v
    d, ha, _, _ = sets
    p = str(d / ("dmg_%s.vskidx" % damage))
    v.index_cache_build(ha, p)
    raw = bytearray(open(p, "rb").read())
    if damage == "truncate":
        raw = raw[: len(raw) // 2]
    elif damage == "magic":
        raw[0] ^= 0xFF
    elif damage == "garbage_len":
        raw[32:40] = b"\xff" * 8    # first array length
    else:
        raw = bytearray()
    open(p, "wb").write(bytes(raw))
    assert not v.index_cache_verify(ha, p)
    assert v.index_cache_build(ha, p) is False
    assert v.index_cache_verify(ha, p)


@pytest.mark.parametrize("threads", ["1", "2", "7", "32"])
def test_threaded_build_equals_one_thread_build(tmp_path, monkeypatch, threads):
    """the index is built on several threads (key ranges sorted independently, the reference's own
    sort replayed over the multi-hit buckets in the reference's order); vsk_index_cache_verify rebuilds
    it with the plain one-thread builder and compares every array -- for any thread count, with
    redundancy groups, long patterns (all 14 seed shapes) and low-complexity consensus strings"""
    import vntrseek_b200 as v
    from vntrseek_b200 import synth
    monkeypatch.setenv("VSK_INDEX_THREADS", threads)
    refs = synth.fast_refs(6000, seed=int(threads) + 3, rot_frac=0.03)
    ts = v.TRSetHandle.from_packed(refs, is_read=False)
    p = str(tmp_path / "t.vskidx")
    assert v.index_cache_build(ts, p) is False
    assert v.index_cache_verify(ts, p)
    monkeypatch.setenv("VSK_INDEX_SERIAL", "1")     # and the switch that selects the one-thread build everywhere
    q = str(tmp_path / "s.vskidx")
    assert v.index_cache_build(ts, q) is False
    assert open(p, "rb").read() == open(q, "rb").read()
    ts.close()
