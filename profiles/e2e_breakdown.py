"""Where the end-to-end time of bench.py goes (single GPU, config 3): model
construction + IC upload, first step, K-1 steps, read-back into pinned memory."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]
import bench
import finite_difference as fd
shape = (512, 512, 512)
pinned = torch.empty(512 ** 3, dtype=torch.float64, pin_memory=True)
ic = pinned.numpy().reshape(shape); ic[...] = 0.5
res = torch.empty(pinned.shape, dtype=pinned.dtype, pin_memory=True).numpy().reshape(shape)
def t(): torch.cuda.synchronize(); return time.perf_counter()
K = 100
for rep in range(3):
    t0 = t(); m, F, step = bench.build_model(fd, "c3", K + 1, ic); t1 = t()
    step(); t2 = t()
    for _ in range(K - 1):
        step()
    t3 = t()
    out = F.local_level(K, out=res); t4 = t()
    m.synchronize(); t5 = t()
    print(f"rep {rep}: build+upload {1e3*(t1-t0):.1f} ms, first step {1e3*(t2-t1):.1f}, {K-1} steps {1e3*(t3-t2):.1f}, "
          f"read-back {1e3*(t4-t3):.1f}, sync {1e3*(t5-t4):.1f}, total {1e3*(t5-t0):.1f}")
    del m, F, step
# raw copies for comparison
d = torch.empty(512 ** 3, dtype=torch.float64, device="cuda")
for _ in range(2):
    a = t(); d.copy_(pinned, non_blocking=True); b = t(); pinned.copy_(d, non_blocking=True); c = t()
print(f"raw 1 GiB H2D {1e3*(b-a):.1f} ms, D2H {1e3*(c-b):.1f} ms")
