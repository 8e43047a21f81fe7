"""GPU: size-independent properties of the PROCLU pipeline on a workload-shaped set
(no reference binary involved): determinism, invariance to how the reads are chunked inside
one call (key budget) and to how they are sharded across calls/ranks, counters consistent
with the link table, and the clustering of the link table independent of shard order."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def workload(ctx):
    import vntrseek_b200 as v
    from vntrseek_b200 import synth
    refs = synth.fast_refs(15000, seed=41)
    reads = synth.fast_reads(refs, 30000, seed=42, rot_frac=0.0)
    rt = v.TRSetHandle.from_packed(refs, is_read=False)
    idx = v.RefIndex(ctx, rt)
    yield refs, reads, idx
    idx.close()
    rt.close()


def _run(ctx, idx, packed, **kw):
    import vntrseek_b200 as v
    ts = v.TRSetHandle.from_packed(packed, is_read=True)
    links, st = v.proclu(ctx, idx, ts, **kw)
    ts.close()
    return links, st


def test_deterministic_and_counters(ctx, workload):
    refs, reads, idx = workload
    a, sa = _run(ctx, idx, reads)
    b, sb = _run(ctx, idx, reads)
    assert len(a) > 1000 and (a == b).all()
    keys = ("candidates_considered", "lcs_alignments", "profile_alignments", "profile_alignments_success", "confirmed")
    assert [sa[k] for k in keys] == [sb[k] for k in keys]
    assert sa["confirmed"] == len(a)
    assert sa["candidates_considered"] >= sa["lcs_alignments"] >= sa["profile_alignments"] >= sa["profile_alignments_success"]
    # every link passed the flank thresholds and the similarity cutoff
    sim = 1.0 - a["dsum"] / (255.0 * a["sumw"])
    assert (sim >= 0.88).all() and (a["lerr"] <= 8).all() and (a["rerr"] <= 8).all()
    # links come out grouped by read, in read order (the order ClusterBaseAddLink would see)
    first = {}
    for i, k in enumerate(a["readkey"]):
        first.setdefault(int(k), i)
    order = sorted(first, key=first.get)
    assert order == sorted(order)


def test_chunking_invariance(ctx, workload):
    refs, reads, idx = workload
    a, sa = _run(ctx, idx, reads)
    b, sb = _run(ctx, idx, reads, key_budget=200000)   # forces many chunks
    assert sb["chunks"] > 5 * sa["chunks"]
    assert (a == b).all()
    for k in ("candidates_considered", "lcs_alignments", "profile_alignments", "profile_alignments_success", "confirmed"):
        assert sa[k] == sb[k], k


def test_sharding_invariance_and_clustering(ctx, workload, tmp_path):
    import vntrseek_b200 as v
    from vntrseek_b200.dist import shard_bounds
    refs, reads, idx = workload
    whole, sw = _run(ctx, idx, reads)
    parts, cand = [], 0
    for r in range(3):
        lo, hi = shard_bounds(len(reads), r, 3)
        l, s = _run(ctx, idx, reads.slice(lo, hi))
        parts.append(l)
        cand += s["candidates_considered"]
    cat = np.concatenate(parts)
    assert (cat == whole).all() and cand == sw["candidates_considered"]
    # cluster membership does not depend on the order in which shards are merged
    def clusters(links, name):
        p = str(tmp_path / name)
        v.write_clusters(links, p)
        out = set()
        for line in open(p):
            out.add(frozenset(int(t[:-1]) for t in line.split()[1:]))
        return out
    assert clusters(whole, "a.clu") == clusters(np.concatenate(parts[::-1]), "b.clu")


def test_links_left_on_device_equal_host_links(ctx, workload):
    """vsk_proclu_result_links_dev: the table the multi-GPU merge sends is the table the host gets, also
    when the run is cut into chunks (the chunks append to the device copy); viewed as a torch tensor
    without a copy (what dist.LinkGatherer takes)"""
    import torch

    import vntrseek_b200 as v
    from vntrseek_b200._lib import LINK_DTYPE
    refs, reads, idx = workload
    ts = v.TRSetHandle.from_packed(reads, is_read=True)
    host, st = v.proclu(ctx, idx, ts, key_budget=200000)
    host = np.array(host)
    dl, st2 = v.proclu(ctx, idx, ts, key_budget=200000, links_on_device=True)
    assert len(dl) == len(host) == st2["confirmed"] and st2["chunks"] > 3
    t = dl.tensor()
    assert t.is_cuda and t.dtype == torch.int32 and tuple(t.shape) == (len(host), 7)
    assert t.data_ptr() == dl.ptr      # a view of the library's buffer, not a copy
    got = t.cpu().numpy().reshape(-1).view(LINK_DTYPE)
    assert (got == host).all()
    ts.close()


def test_gpu_decoded_read_file_equals_text_parsed(ctx, workload, tmp_path):
    """vsk_trset_load_leb36_raw: the read file only tokenised on the host, LEB36 digits / consensus / flank
    clean-up done by the GPU at upload -- same links, same counters as the host parser's set (mixed-case
    and non-ACGT flank characters, dash-majority columns, rotindex included)"""
    import vntrseek_b200 as v
    refs, reads, idx = workload
    path = str(tmp_path / "reads.leb36")
    sub = reads.slice(0, 6000)
    sub.write_leb36(path)
    # spoil some flanks: lower case and non-ACGT characters must come out as the host parser has them
    lines = open(path).read().splitlines()
    for k in range(0, len(lines), 7):
        f = lines[k].rsplit(" ", 1)
        fl = f[1]
        lines[k] = f[0] + " " + fl[: len(fl) // 2].lower() + fl[len(fl) // 2:].replace("A", "R", 1)
    open(path, "w").write("\n".join(lines) + "\n")
    a = v.TRSetHandle.from_file(path, is_read=True, maxflank=50)
    b = v.TRSetHandle.from_file(path, is_read=True, maxflank=50, raw=True)
    la, sa = v.proclu(ctx, idx, a)
    lb, sb = v.proclu(ctx, idx, b)
    keys = ("candidates_considered", "lcs_alignments", "profile_alignments", "profile_alignments_success", "confirmed")
    assert [sa[k] for k in keys] == [sb[k] for k in keys]
    assert len(la) > 500 and (np.array(la) == np.array(lb)).all()
    # and resident, twice (the host learnt the consensus lengths at the first upload)
    rb = v.ResidentReads(ctx, b)
    lc, sc = v.proclu(ctx, idx, rb)
    assert (np.array(lc) == np.array(la)).all()
    rb.close()
    a.close()
    b.close()
