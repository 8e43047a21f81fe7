"""Multi-GPU plumbing: the PROCLU path shards by read TRs (reference: one psearch
process per read file, perl/run_proclu.pl:95-125) and has exactly one exchange
step, the merge of the per-rank link tables (reference: join_clusters over the
per-file .clu outputs, src/join_clusters/joinc.c:246-317).  One process per GPU,
torch.distributed for the collective (NCCL on GPUs, gloo in the CPU tests).

On GPUs the table never passes through the host on its way into the collective: the
library leaves it in HBM (vsk_proclu_result_links_dev), the ranks exchange their record
counts, one all_gather_into_tensor moves the records over NVLink, and only the rank that
writes the cluster file copies the merged table to (page-locked) host memory."""
import numpy as np

from ._lib import LINK_DTYPE

_NF = LINK_DTYPE.itemsize // 4


def shard_bounds(n, rank, world):
    """contiguous range of read TRs owned by `rank`"""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class LinkGatherer:
    """All-gather of per-rank link tables that are resident in HBM.  Buffers (send staging,
    receive, page-locked host copy on the root) are kept between calls and grow on demand."""

    def __init__(self, device, root=0):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.device = torch.device(device)
        self.root = root
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.cap = 0
        self.send = self.recv = self.host = None
        self.counts = torch.zeros(self.world, dtype=torch.int64, device=self.device)
        self.phase_ms = {}

    def _reserve(self, mx):
        if mx <= self.cap:
            return
        torch = self.torch
        self.cap = int(mx * 1.25) + 1024
        self.send = torch.zeros((self.cap, _NF), dtype=torch.int32, device=self.device)
        self.recv = torch.empty((self.world, self.cap, _NF), dtype=torch.int32, device=self.device)
        if self.rank == self.root:
            self.host = torch.empty((self.world, self.cap, _NF), dtype=torch.int32,
                                    pin_memory=self.device.type == "cuda")

    def gather(self, links_dev, to_host=True):
        """links_dev: int32 tensor [n, 7] on this rank's GPU (api.DeviceLinks.tensor()).
        Returns, on the root, the merged table as a structured array in rank order (a view of the
        page-locked buffer: consume it before the next call); None on the other ranks.  With
        to_host=False every rank gets (recv tensor [world, cap, 7], counts list) and nothing is copied.
        self.phase_ms accumulates host wall time per phase (counts exchange incl. the wait for the slowest
        rank, staging copy, all-gather, host copy)."""
        import time
        torch, dist = self.torch, self.dist
        t0 = time.perf_counter()
        n = int(links_dev.shape[0])
        mine = torch.tensor([n], dtype=torch.int64).to(self.device, non_blocking=True)
        dist.all_gather_into_tensor(self.counts, mine)
        counts = [int(c) for c in self.counts.tolist()]   # waits for every rank to have arrived
        t1 = time.perf_counter()
        self._reserve(max(max(counts), 1))
        mx = max(max(counts), 1)
        if n:
            self.send[:n].copy_(links_dev)
        # every rank sends `mx` records (the tail past its count is ignored by the receiver)
        recv = self.recv.view(-1)[: self.world * mx * _NF].view(self.world, mx, _NF)
        dist.all_gather_into_tensor(recv.view(-1), self.send[:mx].view(-1))
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()
        t2 = time.perf_counter()
        self.phase_ms["counts+wait"] = self.phase_ms.get("counts+wait", 0.0) + (t1 - t0) * 1e3
        self.phase_ms["allgather"] = self.phase_ms.get("allgather", 0.0) + (t2 - t1) * 1e3
        if not to_host:
            return recv, counts
        if self.rank != self.root:
            return None
        host = self.host.view(-1)[: self.world * mx * _NF].view(self.world, mx, _NF)
        for r in range(self.world):   # contiguous pieces: plain pinned copies
            if counts[r]:
                host[r, :counts[r]].copy_(recv[r, :counts[r]], non_blocking=True)
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()
        parts = [host[r, :counts[r]].numpy() for r in range(self.world)]
        out = np.concatenate(parts).reshape(-1).view(LINK_DTYPE) if sum(counts) else np.zeros(0, LINK_DTYPE)
        self.phase_ms["host_copy"] = self.phase_ms.get("host_copy", 0.0) + (time.perf_counter() - t2) * 1e3
        return out


def allgather_links(links, device=None):
    """Host-side variant (gloo / CPU tests, or link tables that already live in host memory):
    every rank's link records, concatenated in rank order (all ranks get the result).
    `links`: structured array (LINK_DTYPE).  Uses the default process group; a single
    process returns its input."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return links
    world = dist.get_world_size()
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    n = torch.tensor([len(links)], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    mx = max(max(counts), 1)
    buf = torch.zeros((mx, _NF), dtype=torch.int32, device=dev)
    if len(links):
        buf[:len(links)] = torch.from_numpy(np.ascontiguousarray(links).view(np.int32).reshape(-1, _NF)).to(dev)
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    parts = [out[r][:counts[r]].cpu().numpy() for r in range(world)]
    return np.concatenate(parts).reshape(-1).view(LINK_DTYPE) if sum(counts) else np.zeros(0, LINK_DTYPE)
