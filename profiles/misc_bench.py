"""Step times of a few everyday model shapes (not BASELINE configs): looks for
performance cliffs next to the benchmarked cases."""
import contextlib, io, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]
import torch
import finite_difference as fd


def heat(shape, periodic, dx=1.0, steps=30, form="k*(sum)", radius=1):
    names = ["x", "y", "z"][:len(shape)]
    axes = {"t": np.arange(steps + 8) * 1e-3}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * dx
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=tuple(names) if periodic else (), history="last")
        T = fd.Field(m, "T", n_time_ders=1)
        st = fd.Stencil(list(range(-radius, radius + 1)), 2)
        T.set_IC_values(np.random.default_rng(0).random(shape))
        if not periodic:
            for a in names:
                T.set_BC("0", a, "start")
                T.set_BC("1", a, "end")

    def step():
        terms = [st.der(T.prev, a) for a in names]
        if form == "k*(sum)":
            acc = terms[0]
            for tm in terms[1:]:
                acc = acc + tm
            expr = 0.1 * acc
        else:
            acc = 0.1 * terms[0]
            for tm in terms[1:]:
                acc = acc + 0.1 * tm
            expr = acc
        T.dot.assign_update(expr)
        T.time_integrate_update()
        m.increment_time()
    for _ in range(5):
        step()
    m.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    pts = float(np.prod(shape))
    return ms, pts * 16 / (ms * 1e-3) / 1e9, m._engine.stats


CASES = [
    ("1-D 16M periodic", (1 << 24,), True, 1.0, "k*(sum)", 1),
    ("1-D 16M fixed", (1 << 24,), False, 1.0, "k*(sum)", 1),
    ("2-D 4096^2 periodic", (4096, 4096), True, 1.0, "k*(sum)", 1),
    ("2-D 4000x4100 fixed", (4000, 4100), False, 0.01, "k*(sum)", 1),
    ("3-D 500^3 periodic", (500, 500, 500), True, 1.0, "k*(sum)", 1),
    ("3-D 512^3 fixed", (512, 512, 512), False, 1.0, "k*(sum)", 1),
    ("3-D 512^3 fixed dx=0.01", (512, 512, 512), False, 0.01, "k*(sum)", 1),
    ("3-D 512^3 periodic k*dx+k*dy+k*dz", (512, 512, 512), True, 1.0, "sum of k*", 1),
    ("3-D 512^3 periodic radius 2", (512, 512, 512), True, 1.0, "k*(sum)", 2),
    ("3-D 384^3 periodic radius 4", (384, 384, 384), True, 1.0, "k*(sum)", 4),
    ("3-D 64^3 periodic (launch-bound)", (64, 64, 64), True, 1.0, "k*(sum)", 1),
]
import os as _os
_only = _os.environ.get("MISC_ONLY")
for name, shape, periodic, dx, form, radius in CASES:
    if _only and _only not in name:
        continue
    ms, gbs, stats = heat(shape, periodic, dx, form=form, radius=radius)
    print(f"{name:40s} {ms:8.4f} ms/step  {gbs:7.0f} GB/s algorithmic  {stats}")
