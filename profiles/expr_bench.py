"""Step times of statements that take the postfix-program kernel (fd_fused_expr) and the nested-der path
(fd_fused_nested): Burgers 2-D at 8192^2, a reaction-diffusion term at 512^3, div(kappa grad T) at 4096^2.
Algorithmic bytes: one read + one write of every integrated level (16 B per point in fp64); the nested case also
reads the coefficient field and writes + reads its two scratch levels (56 B per point)."""
import contextlib, io, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]
import torch
import finite_difference as fd


def timed(m, step, steps=20, warm=4):
    for _ in range(warm):
        step()
    m.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def burgers2d(n=8192, steps=20):
    axes = {"t": np.arange(steps + 8) * 1e-4, "x": np.arange(n) * 1.0, "y": np.arange(n) * 1.0}
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=("x", "y"), history="last")
        u = fd.Field(m, "u", n_time_ders=1)
        d1, d2 = fd.Stencil([-1, 0, 1], 1), fd.Stencil([-1, 0, 1], 2)
        u.set_IC_values(1.0 + np.random.default_rng(0).random((n, n)))

    def step():
        adv = u.prev * d1.der(u.prev, "x") + u.prev * d1.der(u.prev, "y")
        u.dot.assign_update(0.05 * (d2.der(u.prev, "x") + d2.der(u.prev, "y")) - adv)
        u.time_integrate_update()
        m.increment_time()
    return m, step, float(n) * n, 16


def reaction3d(n=512, steps=20):
    axes = {"t": np.arange(steps + 8) * 1e-4}
    for a in "xyz":
        axes[a] = np.arange(n) * 1.0
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=("x", "y", "z"), history="last")
        u = fd.Field(m, "u", n_time_ders=1)
        d2 = fd.Stencil([-1, 0, 1], 2)
        u.set_IC_values(np.random.default_rng(0).random((n, n, n)))

    def step():
        lap = d2.der(u.prev, "x") + d2.der(u.prev, "y") + d2.der(u.prev, "z")
        u.dot.assign_update(0.1 * lap + u.prev * (1 - u.prev ** 2))
        u.time_integrate_update()
        m.increment_time()
    return m, step, float(n) ** 3, 16


def divgrad2d(n=4096, steps=20):
    axes = {"t": np.arange(steps + 8) * 1e-4, "x": np.arange(n) * 1.0, "y": np.arange(n) * 1.0}
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=("x", "y"), history="last")
        T = fd.Field(m, "T", n_time_ders=1)
        kappa = fd.ConstantField(m, "kappa")
        kappa.set("1+x/8192+y/8192")
        d1 = fd.Stencil([-1, 0, 1], 1)
        T.set_IC_values(np.random.default_rng(0).random((n, n)))

    def step():
        T.dot.assign_update(d1.der(kappa.always * d1.der(T.prev, "x"), "x") + d1.der(kappa.always * d1.der(T.prev, "y"), "y"))
        T.time_integrate_update()
        m.increment_time()
    return m, step, float(n) * n, 56


only = os.environ.get("EXPR_ONLY")
peak = 6543.1
for name, make in (("burgers 2-D 8192^2 (expr kernel)", burgers2d), ("reaction-diffusion 512^3 (expr kernel)", reaction3d),
                   ("div(kappa grad T) 4096^2 (nested: 2 scratch levels + statement)", divgrad2d)):
    if only and only not in name:
        continue
    m, step, pts, bpp = make()
    ms = timed(m, step)
    gbs = pts * bpp / (ms * 1e-3) / 1e9
    print(f"{name:66s} {ms:8.4f} ms/step  {pts / (ms * 1e-3):.3e} pt/s  {gbs:6.0f} GB/s algorithmic = {gbs / peak:.2f} of the measured HBM peak  {m._engine.stats}")
    m.close()
