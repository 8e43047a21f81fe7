"""all_gather_into_tensor timing probe (torchrun, N ranks): bytes per rank vs time, NCCL transport from NCCL_DEBUG=INFO"""
import os
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for mb in (1, 8, 47, 188):
    n = mb * 1024 * 1024 // 4
    send = torch.ones(n, dtype=torch.int32, device="cuda")
    recv = torch.empty(n * world, dtype=torch.int32, device="cuda")
    for _ in range(3):
        dist.all_gather_into_tensor(recv, send)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        dist.all_gather_into_tensor(recv, send)
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / 10
    host = torch.empty(n * world, dtype=torch.int32, pin_memory=True)
    e0.record()
    host.copy_(recv, non_blocking=True)
    e1.record()
    e1.synchronize()
    if rank == 0:
        print("all_gather %4d MB/rank x %d ranks: %.3f ms (%.1f GB/s per rank in); D2H of the gathered %d MB: %.2f ms"
              % (mb, world, ms, mb * (world - 1) / 1024 / (ms * 1e-3), mb * world, e0.elapsed_time(e1)), flush=True)
if rank == 0:
    print("can_device_access_peer(0,1):", torch.cuda.can_device_access_peer(0, 1))
dist.destroy_process_group()
