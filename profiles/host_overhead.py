"""Host-side cost of one recorded time step (tiny grid, so the GPU is idle):
the Python layer must stay well below the GPU step time of the big grids."""
import contextlib, io, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]
import bench
import finite_difference as fd
import torch

bench.WORKLOADS["tiny"] = ("tiny", (32, 32, 32), 16)
ic = np.random.default_rng(0).random((32, 32, 32))
os.environ["FD_B200_POLL_EVERY"] = "64"
shape_bak = bench.WORKLOADS["c3"]
bench.WORKLOADS["c3"] = ("tiny", (32, 32, 32), 16)
m, T, step = bench.build_model(fd, "c3", 2100, ic)
for _ in range(50):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(2000):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host cost per step: {(t1 - t0) / 2000 * 1e6:.1f} us (+{(t2 - t1) * 1e3:.2f} ms drain)")
