"""Host<->device copy rates of 1 GiB pinned buffers when one rank copies alone and when all ranks copy at once
(torchrun --nproc-per-node N profiles/pcie_probe.py): names the limiter of the end-to-end figure on several GPUs,
whose timed region uploads every rank's initial level and reads its final level back."""
import os, time
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 27
host = torch.empty(n, dtype=torch.float64, pin_memory=True)
host.fill_(1.0)
dev = torch.empty(n, dtype=torch.float64, device="cuda")


def timed(direction):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if direction == "h2d":
        dev.copy_(host, non_blocking=True)
    else:
        host.copy_(dev, non_blocking=True)
    torch.cuda.synchronize()
    return n * 8 / (time.perf_counter() - t0) / 1e9


for direction in ("h2d", "d2h"):
    timed(direction)
    alone = torch.zeros(world, dtype=torch.float64, device="cuda")
    for r in range(world):                      # one rank at a time
        dist.barrier()
        if r == rank:
            alone[r] = timed(direction)
    dist.barrier()
    together = torch.zeros(world, dtype=torch.float64, device="cuda")
    together[rank] = timed(direction)           # everybody at once
    dist.all_reduce(alone)
    dist.all_reduce(together)
    if rank == 0:
        print(f"{direction} alone    GB/s per rank:", " ".join(f"{v:5.1f}" for v in alone.tolist()))
        print(f"{direction} together GB/s per rank:", " ".join(f"{v:5.1f}" for v in together.tolist()),
              f"  sum {together.sum().item():6.1f}")
dist.destroy_process_group()
