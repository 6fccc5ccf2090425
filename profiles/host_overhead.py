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


# the same steps replayed by Model.run (the body is called twice, then its statements are re-issued)
import contextlib, io
with contextlib.redirect_stdout(io.StringIO()):
    m2 = fd.Model({"t": np.arange(2100) * 1e-3, "x": np.arange(32) * 1.0, "y": np.arange(32) * 1.0, "z": np.arange(32) * 1.0},
                  "t", periodic=("x", "y", "z"), history="last")
    T2 = fd.Field(m2, "T", n_time_ders=1)
    st2 = fd.Stencil([-1, 0, 1], 2)
    T2.set_IC_values(ic)


def body2():
    T2.dot.assign_update(0.1 * (st2.der(T2.prev, "x") + st2.der(T2.prev, "y") + st2.der(T2.prev, "z")))
    T2.time_integrate_update()


m2.run(body2, 50)
torch.cuda.synchronize()
t0 = time.perf_counter()
m2.run(body2, 2000)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"Model.run: host cost per step {(t1 - t0) / 2000 * 1e6:.1f} us (+{(t2 - t1) * 1e3:.2f} ms drain)")
