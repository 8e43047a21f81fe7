"""CPU, world_size 2 over gloo: the N>1 host logic -- read sharding and the link-table
all-gather -- gives every rank the table a single process would have produced."""
import os
import socket

import numpy as np
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _links(n0, n1, seed):
    from vntrseek_b200._lib import LINK_DTYPE
    rng = np.random.default_rng(seed)
    a = np.zeros(n1 - n0, LINK_DTYPE)
    a["readkey"] = np.arange(n0, n1) + 1
    a["refkey"] = -rng.integers(1, 50, n1 - n0)
    a["dir"] = rng.integers(0, 2, n1 - n0)
    a["lerr"] = rng.integers(0, 9, n1 - n0)
    a["rerr"] = rng.integers(0, 9, n1 - n0)
    a["dsum"] = rng.integers(0, 1000, n1 - n0)
    a["sumw"] = rng.integers(1, 1000, n1 - n0)
    return a


def _worker(rank, world, port, n, outdir):
    import torch.distributed as dist

    from vntrseek_b200.dist import allgather_links, shard_bounds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_bounds(n, rank, world)
    full = _links(0, n, 7)
    mine = full[lo:hi] if rank != 1 else full[lo:hi][:0] if n == 3 else full[lo:hi]
    got = allgather_links(mine)
    np.save(os.path.join(outdir, "r%d.npy" % rank), got)
    dist.destroy_process_group()


def _worker_gatherer(rank, world, port, n, outdir):
    import torch
    import torch.distributed as dist

    from vntrseek_b200.dist import LinkGatherer, shard_bounds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = LinkGatherer("cpu", root=0)
    full = _links(0, n, 7)
    for rep, cut in enumerate((n, n // 3, n)):  # the buffers are re-used: a smaller table after a larger one
        lo, hi = shard_bounds(cut, rank, world)
        mine = full[lo:hi] if not (rank == 1 and cut == 1) else full[:0]
        t = torch.from_numpy(np.ascontiguousarray(mine).view(np.int32).reshape(-1, 7).copy())
        got = g.gather(t)
        assert (got is None) == (rank != 0)
        if rank == 0:
            np.save(os.path.join(outdir, "g%d.npy" % rep), got.wait())
    recv, counts = g.gather(t, to_host=False)
    assert sum(counts) == n and recv.shape[0] == world
    dist.destroy_process_group()


def test_link_gatherer_two_ranks(tmp_path):
    """the device-resident merge (dist.LinkGatherer) run over gloo with host tensors: rank order,
    ragged counts, buffer re-use; only the root receives the host copy"""
    n = 2003
    mp.spawn(_worker_gatherer, args=(2, _free_port(), n, str(tmp_path)), nprocs=2, join=True)
    full = _links(0, n, 7)
    for rep, cut in enumerate((n, n // 3, n)):
        got = np.load(tmp_path / ("g%d.npy" % rep))
        assert got.dtype == full.dtype and (got == full[:cut]).all()


def test_shard_bounds_cover():
    from vntrseek_b200.dist import shard_bounds
    for n in (0, 1, 7, 1000001):
        for w in (1, 2, 3, 8):
            b = [shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


def test_allgather_links_two_ranks(tmp_path):
    n = 1001
    mp.spawn(_worker, args=(2, _free_port(), n, str(tmp_path)), nprocs=2, join=True)
    full = _links(0, n, 7)
    for r in range(2):
        got = np.load(tmp_path / ("r%d.npy" % r))
        assert got.dtype == full.dtype and len(got) == n
        assert (got == full).all()


def test_allgather_links_empty_rank(tmp_path):
    # n == 3: rank 1 contributes nothing (ragged / empty input)
    mp.spawn(_worker, args=(2, _free_port(), 3, str(tmp_path)), nprocs=2, join=True)
    from vntrseek_b200.dist import shard_bounds
    full = _links(0, 3, 7)
    lo, hi = shard_bounds(3, 0, 2)
    for r in range(2):
        got = np.load(tmp_path / ("r%d.npy" % r))
        assert (got == full[lo:hi]).all()
