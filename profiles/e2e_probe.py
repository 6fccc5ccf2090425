import os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path[:0] = ["/root/repo", "/root/repo/finite-difference_b200"]
os.chdir("/root/repo")
import bench
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=bench._nccl_options())
import finite_difference as fd
rank, world = dist.get_rank(), dist.get_world_size()
shape = (512, 512, 512)
pinned = torch.empty(512**3, dtype=torch.float64, pin_memory=True)
ic = pinned.numpy().reshape(shape); ic[...] = 0.5
def t(): torch.cuda.synchronize(); return time.perf_counter()
for rep in range(2):
    t0 = t(); m, F, step = bench.build_model(fd, "c3", 21, ic, world=world); t1 = t()
    step(); t2 = t()
    per = []
    for _ in range(19):
        a = t(); step(); per.append(t() - a)
    t3 = t()
    if rank == 0: print("per-step ms:", " ".join(f"{x*1e3:.1f}" for x in per))
    out = F.local_level(20); t4 = t()
    m.synchronize(); t5 = t()
    if rank == 0: print(f"rep {rep}: build {t1-t0:.3f} first step {t2-t1:.3f} 19 steps {t3-t2:.3f} local_level {t4-t3:.3f} sync {t5-t4:.3f}")
    del m, F, step
dist.destroy_process_group()
