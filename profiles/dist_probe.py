"""Where does a multi-GPU step spend its time?  Run under torchrun on >= 2 GPUs:
times (1) the halo exchange alone, (2) the step with the exchange disabled
(results wrong, timing only), (3) the full step; host-side and device-side."""
import os, sys, time
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]


def _nccl_options():
    """NCCL kernels on a high-priority stream, so that a halo exchange issued
    while the interior kernel still has blocks to dispatch is not queued behind it."""
    import torch.distributed as dist
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = True
    return opts

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=_nccl_options())
import bench
import finite_difference as fd
from finite_difference import _dist
rank, world = dist.get_rank(), dist.get_world_size()
shape = (512, 512, 512)
ic = np.random.default_rng(rank).random(shape)
m, T, step = bench.build_model(fd, "c3", 400, ic, world=world)
for _ in range(5):
    step()
torch.cuda.synchronize(); dist.barrier()

def timed(fn, n):
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); host = time.perf_counter() - t0
    torch.cuda.synchronize()
    return host / n * 1e6, e0.elapsed_time(e1) / n * 1e3

st = m._engine.state(T)
def exch():
    m._engine.slab.exchange(st.ring[0], st.plane, st.owned0, st.H)
h, d = timed(exch, 100)
if rank == 0: print(f"exchange alone      : host {h:7.1f} us  device {d:7.1f} us")
h, d = timed(step, 100)
if rank == 0: print(f"full step           : host {h:7.1f} us  device {d:7.1f} us")
real = _dist.Slab.exchange
_dist.Slab.exchange = lambda self, *a: None
h, d = timed(step, 100)
if rank == 0: print(f"step, exchange off  : host {h:7.1f} us  device {d:7.1f} us")
_dist.Slab.exchange = real
dist.barrier(); dist.destroy_process_group()
