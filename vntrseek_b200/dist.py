"""Multi-GPU plumbing: the PROCLU path shards by read TRs (reference: one psearch
process per read file, perl/run_proclu.pl:95-125) and has exactly one exchange
step, the merge of the per-rank link tables (reference: join_clusters over the
per-file .clu outputs, src/join_clusters/joinc.c:246-317).  One process per GPU,
torch.distributed for the collective (NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np

from ._lib import LINK_DTYPE


def shard_bounds(n, rank, world):
    """contiguous range of read TRs owned by `rank`"""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allgather_links(links, device=None):
    """every rank's link records, concatenated in rank order (all ranks get the result).
    `links`: structured array (LINK_DTYPE).  Uses the default process group; a single
    process returns its input."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return links
    world = dist.get_world_size()
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    nf = LINK_DTYPE.itemsize // 4
    n = torch.tensor([len(links)], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    mx = max(max(counts), 1)
    buf = torch.zeros((mx, nf), dtype=torch.int32, device=dev)
    if len(links):
        buf[:len(links)] = torch.from_numpy(np.ascontiguousarray(links).view(np.int32).reshape(-1, nf)).to(dev)
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    parts = [out[r][:counts[r]].cpu().numpy() for r in range(world)]
    return np.concatenate(parts).reshape(-1).view(LINK_DTYPE) if sum(counts) else np.zeros(0, LINK_DTYPE)
