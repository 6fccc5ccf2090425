"""Why is config 2's host cost per step higher inside bench.py's workloads block than in a dedicated run?
Measures the host time of 20 c2 steps (a) first thing in the process, (b) after a c3 model has run and been closed,
(c) after a history='all' run, with a cProfile of (c)."""
import cProfile, io, os, pstats, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]
import torch
import bench
import finite_difference as fd


def c2_host(tag, profile=False):
    shape = bench.WORKLOADS["c2"][1]
    ic = np.random.default_rng(0).random(shape)
    m, F, step = bench.build_model(fd, "c2", 40, ic)
    for _ in range(5):
        step()
    m.synchronize()
    torch.cuda.synchronize()
    pr = cProfile.Profile() if profile else None
    t0 = time.perf_counter()
    if pr:
        pr.enable()
    for _ in range(20):
        step()
    if pr:
        pr.disable()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"{tag:32s} host {1e3 * (t1 - t0) / 20:.4f} ms/step, with drain {1e3 * (t2 - t0) / 20:.4f} ms/step, stats {m._engine.stats}, seq {m._engine.seq}")
    if pr:
        s = io.StringIO()
        pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(14)
        print(s.getvalue()[:2600])
    m.close()


c2_host("fresh process")
c2_host("second c2 model")
ic3 = np.random.default_rng(0).random((512, 512, 512))
m, F, step = bench.build_model(fd, "c3", 40, ic3)
for _ in range(25):
    step()
m.synchronize()
m.close()
torch.cuda.empty_cache()
c2_host("after a c3 model")
m, F, step = bench.build_model(fd, "c3", 5, ic3, history="all")
for _ in range(4):
    step()
F.local_level(4)
m.synchronize()
m.close()
del m, F, step
torch.cuda.empty_cache()
c2_host("after a history='all' model", profile=True)
import gc
gc.collect()
c2_host("after gc.collect()")
