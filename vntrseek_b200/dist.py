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


_D2H_ROWS = 1 << 18   # 7 MB per piece


class MergedLinks:
    """The merged link table on the root while its copy to page-locked host memory may still be running
    (it overlaps whatever the caller does next, e.g. the next scan).  `.wait()` returns the table as a
    structured array in rank order: a view of the gatherer's host buffer, valid until the gatherer's
    next-but-one call (the buffers alternate)."""

    def __init__(self, host, total, event):
        self._host, self._total, self._event = host, total, event

    def wait(self):
        if self._event is not None:
            self._event.synchronize()
            self._event = None
        if self._total == 0:
            return np.zeros(0, LINK_DTYPE)
        return self._host[: self._total].numpy().reshape(-1).view(LINK_DTYPE)


class LinkGatherer:
    """All-gather of per-rank link tables that are resident in HBM.  Buffers (send staging, receive, packed
    table, two page-locked host copies on the root) are kept between calls and grow on demand."""

    def __init__(self, device, root=0):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.device = torch.device(device)
        self.root = root
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.cap = 0
        self.send = self.recv = self.packed = None
        self.host = [None, None]
        self.flip = 0
        self.counts = torch.zeros(self.world, dtype=torch.int64, device=self.device)
        self.phase_ms = {}
        self.copy_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self.last_copy = None  # event of the previous device-to-host copy (it reads `packed`)

    def _reserve(self, mx):
        if mx <= self.cap:
            return
        torch = self.torch
        if self.last_copy is not None:
            self.last_copy.synchronize()
        self.cap = int(mx * 1.25) + 1024
        self.send = torch.zeros((self.cap, _NF), dtype=torch.int32, device=self.device)
        self.recv = torch.empty((self.world * self.cap, _NF), dtype=torch.int32, device=self.device)
        if self.rank == self.root:
            self.packed = torch.empty((self.world * self.cap, _NF), dtype=torch.int32, device=self.device)
            self.host = [torch.empty((self.world * self.cap, _NF), dtype=torch.int32,
                                     pin_memory=self.device.type == "cuda") for _ in range(2)]

    def gather(self, links_dev, to_host=True):
        """links_dev: int32 tensor [n, 7] on this rank's GPU (api.DeviceLinks.tensor()).
        Returns, on the root, a MergedLinks (`.wait()` gives the merged table in rank order; the copy to the
        host runs on a side stream and overlaps the caller's next work); None on the other ranks.  With
        to_host=False every rank gets (recv tensor [world, max count, 7], counts list) and nothing is copied.
        self.phase_ms accumulates host wall time per phase (counts exchange incl. the wait for the slowest
        rank; staging copy + all-gather)."""
        import time
        torch, dist = self.torch, self.dist
        t0 = time.perf_counter()
        n = int(links_dev.shape[0])
        mine = torch.tensor([n], dtype=torch.int64).to(self.device, non_blocking=True)
        dist.all_gather_into_tensor(self.counts, mine)
        counts = [int(c) for c in self.counts.tolist()]   # waits for every rank to have arrived
        t1 = time.perf_counter()
        mx = max(max(counts), 1)
        self._reserve(mx)
        if n:
            self.send[:n].copy_(links_dev)
        # every rank sends `mx` records (the tail past its count is ignored by the receiver)
        recv = self.recv[: self.world * mx].view(self.world, mx, _NF)
        dist.all_gather_into_tensor(recv.view(-1), self.send[:mx].view(-1))
        out = None
        if to_host and self.rank == self.root:
            # ragged parts packed on the device, then ONE copy into page-locked memory on a side stream
            if self.last_copy is not None:
                self.last_copy.synchronize()     # the previous copy has read `packed` by now
            total, at = sum(counts), 0
            for r in range(self.world):
                if counts[r]:
                    self.packed[at:at + counts[r]].copy_(recv[r, :counts[r]])
                    at += counts[r]
            host = self.host[self.flip]
            self.flip ^= 1
            if self.device.type == "cuda":
                ready = torch.cuda.Event()
                ready.record(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(self.copy_stream):
                    self.copy_stream.wait_event(ready)
                    # in pieces, so that the small copies of the caller's next scan are not queued behind
                    # one 300 MB transfer on the copy engine
                    for a in range(0, total, _D2H_ROWS):
                        b = min(total, a + _D2H_ROWS)
                        host[a:b].copy_(self.packed[a:b], non_blocking=True)
                    done = torch.cuda.Event()
                    done.record(self.copy_stream)
                self.last_copy = done
                out = MergedLinks(host, total, done)
            else:
                if total:
                    host[:total].copy_(self.packed[:total])
                out = MergedLinks(host, total, None)
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()
        t2 = time.perf_counter()
        self.phase_ms["counts+wait"] = self.phase_ms.get("counts+wait", 0.0) + (t1 - t0) * 1e3
        self.phase_ms["allgather"] = self.phase_ms.get("allgather", 0.0) + (t2 - t1) * 1e3
        if not to_host:
            return recv, counts
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
