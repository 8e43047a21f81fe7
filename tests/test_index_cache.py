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
