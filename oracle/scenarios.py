"""Scenarios written once against the public ``fd`` API.

TEST INFRASTRUCTURE ONLY.  Every function takes the module to drive -- the
unmodified reference (``oracle.ref_loader.load()``, build container only) or
the B200 package (``import finite_difference``) -- and returns a dict of
result arrays.  ``oracle/make_golden.py`` runs them on the reference and
commits the outputs to ``tests/golden/``; the parity tests run them on the
B200 package and compare.

Array initial conditions are injected the way reference users have to do it
(``set_IC`` only takes strings, FD.py:394): write the t=0 slice of
``Field.data`` and set ``IC_set``.
"""

from __future__ import annotations

import numpy as np


def _rng(seed):
    return np.random.default_rng(seed)


def _inject_ic(field, t_index_axis, values):
    """Write ``values`` into the t=0 slice of field.data (time axis position
    ``t_index_axis``)."""
    if getattr(getattr(field, "model", None), "_slab", None) is not None:
        # B200 package on several ranks: Field.data is a gathered copy there
        field.set_IC_values(values)
        return
    sel = [slice(None)] * field.data.ndim
    sel[t_index_axis] = 0
    field.data[tuple(sel)] = values
    field.IC_set = True


def _loop(m, body):
    while not m.finished:
        body()
        m.increment_time()


# -- C.1 ---------------------------------------------------------------------

def readme_1d(fd):
    """README.md:30-75 with the der_axis the README forgets."""
    m = fd.Model({"x": range(1, 100, 4), "t": range(1, 10)}, time_axis="t")
    T = fd.Field(m, "Temperature", n_time_ders=1)
    diff_2 = fd.Stencil([-1, 0, 1], der_order=2)
    T.set_IC("1")
    T.set_BC("0", "x", "start")
    T.set_BC("0", "x", "end")
    k = 2
    m.check_IC()

    def body():
        T.dot.assign_update(k * diff_2.der(T.prev, "x"))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


# -- C.2 ---------------------------------------------------------------------

def heat2d_fixed(fd, nx=12, ny=10, n_t=6, k=0.1, dt=1e-3, seed=0):
    m = fd.Model({"t": np.arange(n_t) * dt, "x": np.arange(nx) * 1.0,
                  "y": np.arange(ny) * 1.0}, "t")
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(T, 0, _rng(seed).random((nx, ny)))
    for a in ("x", "y"):
        T.set_BC("0", a, "start")
        T.set_BC("0", a, "end")

    def body():
        T.dot.assign_update(k * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def heat2d_array_bc(fd, nx=10, ny=9, n_t=5, seed=16):
    """Boundary values written into Field.data by hand for every time level (the
    array counterpart of set_BC, which only takes strings)."""
    m = fd.Model({"t": np.arange(n_t) * 1e-3, "x": np.arange(nx) * 1.0, "y": np.arange(ny) * 1.0}, "t")
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    rng = _rng(seed)
    data = T.data
    data[0] = rng.random((nx, ny))
    data[:, 0, :] = rng.random((n_t, ny))
    data[:, -1, :] = 2.0
    data[:, :, 0] = rng.random((n_t, nx))
    data[:, :, -1] = np.arange(n_t)[:, None] * 0.5
    T.IC_set = True

    def body():
        T.dot.assign_update(0.1 * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


# -- C.3 ---------------------------------------------------------------------

def heat3d_periodic(fd, shape=(6, 5, 4), n_t=7, k=0.1, dt=1e-3, seed=0):
    nx, ny, nz = shape
    m = fd.Model({"t": np.arange(n_t) * dt, "x": np.arange(nx) * 1.0,
                  "y": np.arange(ny) * 1.0, "z": np.arange(nz) * 1.0}, "t",
                 periodic=("x", "y", "z"))
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(T, 0, _rng(seed).random(shape))

    def body():
        lap = d2.der(T.prev, "x") + d2.der(T.prev, "y") + d2.der(T.prev, "z")
        T.dot.assign_update(k * lap)
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def heat3d_fixed(fd, shape=(7, 6, 9), n_t=5, k=0.05, dt=1e-2, seed=1):
    nx, ny, nz = shape
    m = fd.Model({"t": np.arange(n_t) * dt, "x": np.arange(nx) * 0.5,
                  "y": np.arange(ny) * 0.25, "z": np.arange(nz) * 2.0}, "t")
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(T, 0, _rng(seed).random(shape))
    T.set_BC("0", "x", "start")
    T.set_BC("1", "x", "end")
    T.set_BC("z", "y", "start")
    T.set_BC("x+z", "y", "end")
    T.set_BC("t", "z", "start")
    T.set_BC("2*t+x*y", "z", "end")

    def body():
        lap = d2.der(T.prev, "x") + d2.der(T.prev, "y") + d2.der(T.prev, "z")
        T.dot.assign_update(k * lap)
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def heat1d_odd_dx(fd, n=33, n_t=9, seed=2):
    """dx = 0.3 (division by a non-power-of-two), time axis LAST, wide stencil
    on a periodic axis."""
    m = fd.Model({"x": np.arange(n) * 0.3, "t": np.arange(n_t) * 1e-3}, "t",
                 periodic=["x"])
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-2, -1, 0, 1, 2], 2)
    _inject_ic(T, 1, _rng(seed).random(n))

    def body():
        T.dot.assign_update(0.7 * d2.der(T.prev, "x"))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def heat4d_periodic(fd, shape=(4, 3, 5, 6), n_t=4, seed=3):
    names = ("x", "y", "z", "w")
    axes = {"t": np.arange(n_t) * 1e-3}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    m = fd.Model(axes, "t", periodic=names)
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(T, 0, _rng(seed).random(shape))

    def body():
        lap = d2.der(T.prev, "x")
        for a in names[1:]:
            lap = lap + d2.der(T.prev, a)
        T.dot.assign_update(0.1 * lap)
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def heat2d_time_middle(fd, nx=9, ny=8, n_t=5, seed=4):
    """Time axis in the middle of the dict; mixed periodic/fixed axes."""
    m = fd.Model({"x": np.arange(nx) * 1.0, "t": np.arange(n_t) * 1e-2,
                  "y": np.arange(ny) * 1.0}, "t", periodic=("y",))
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(T, 1, _rng(seed).random((nx, ny)))
    T.set_BC("y", "x", "start")
    T.set_BC("1", "x", "end")

    def body():
        T.dot.assign_update(0.2 * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


# -- C.4 ---------------------------------------------------------------------

def wave2d(fd, nx=16, ny=12, n_t=6, c2=0.01, dt=1e-2, points=tuple(range(-4, 5)),
           periodic=True, seed=5):
    m = fd.Model({"t": np.arange(n_t) * dt, "x": np.arange(nx) * 1.0,
                  "y": np.arange(ny) * 1.0}, "t",
                 periodic=("x", "y") if periodic else ())
    u = fd.Field(m, "u", n_time_ders=2)
    d2 = fd.Stencil(list(points), 2)
    _inject_ic(u, 0, _rng(seed).random((nx, ny)))
    _inject_ic(u.dot, 0, np.zeros((nx, ny)))

    def body():
        u.dot.dot.assign_update(c2 * (d2.der(u.prev, "x") + d2.der(u.prev, "y")))
        u.dot.time_integrate_update()
        u.time_integrate_update()

    _loop(m, body)
    return {"u": u.data, "udot": u.dot.data, "udotdot": u.dot.dot.data}


def wave2d_5pt(fd):
    return wave2d(fd, points=(-2, -1, 0, 1, 2), seed=6)


def wave2d_fixed_nan_spread(fd):
    """Fixed axes without BCs: the NaN front advances by the stencil radius
    every step and the 'losing data' note is printed once."""
    return wave2d(fd, nx=24, ny=24, n_t=4, points=(-2, -1, 0, 1, 2), periodic=False, seed=7)


# -- C.5 ---------------------------------------------------------------------

def acoustic3d(fd, shape=(6, 5, 4), n_t=6, rho=1.0, K=0.25, dt=0.1, periodic=True, seed=8):
    names = ("x", "y", "z")
    axes = {"t": np.arange(n_t) * dt}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    m = fd.Model(axes, "t", periodic=names if periodic else ())
    p = fd.Field(m, "p", n_time_ders=1)
    v = {a: fd.Field(m, "v" + a, edge_axes=(a,), n_time_ders=1) for a in names}
    d_ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
    d_ec = fd.Stencil([-.5, .5], 1, "edge", "cell")
    _inject_ic(p, 0, _rng(seed).random(shape))
    for a in names:
        sh = list(shape)
        if not periodic:
            sh[names.index(a)] += 1
        _inject_ic(v[a], 0, np.zeros(sh))
        if not periodic:
            v[a].set_BC("0", a, "start")
            v[a].set_BC("0", a, "end")

    def body():
        for a in names:
            v[a].dot.assign_update((-1 / rho) * d_ce.der(p.prev, a))
            v[a].time_integrate_update()
        div = d_ec.der(v["x"].new, "x") + d_ec.der(v["y"].new, "y") + d_ec.der(v["z"].new, "z")
        p.dot.assign_update((-K) * div)
        p.time_integrate_update()

    _loop(m, body)
    out = {"p": p.data, "pdot": p.dot.data}
    for a in names:
        out["v" + a] = v[a].data
    return out


def acoustic3d_fixed(fd):
    return acoustic3d(fd, shape=(5, 6, 4), periodic=False, seed=9)


def acoustic3d_wide(fd):
    """Periodic staggered step on a grid wide enough for 128-bit vectors along z
    (the B200 build runs it as ONE fused velocity+pressure launch per step)."""
    return acoustic3d(fd, shape=(8, 5, 12), n_t=5, seed=14)


def acoustic2d(fd, shape=(9, 16), n_t=5, seed=15):
    """2-D staggered pair with unequal spacings and a subtraction in the divergence."""
    names = ("x", "y")
    m = fd.Model({"t": np.arange(n_t) * 0.05, "x": np.arange(shape[0]) * 0.5, "y": np.arange(shape[1]) * 2.0},
                 "t", periodic=names)
    p = fd.Field(m, "p", n_time_ders=1)
    v = {a: fd.Field(m, "v" + a, edge_axes=(a,), n_time_ders=1) for a in names}
    d_ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
    d_ec = fd.Stencil([-.5, .5], 1, "edge", "cell")
    _inject_ic(p, 0, _rng(seed).random(shape))
    for a in names:
        _inject_ic(v[a], 0, _rng(seed + 1).random(shape) * 0.1)

    def body():
        for a in names:
            v[a].dot.assign_update(d_ce.der(p.prev, a) * -0.7)
            v[a].time_integrate_update()
        p.dot.assign_update(0.3 * (d_ec.der(v["x"].new, "x") - d_ec.der(v["y"].new, "y")))
        p.time_integrate_update()

    _loop(m, body)
    return {"p": p.data, "vx": v["x"].data, "vy": v["y"].data, "pdot": p.dot.data, "vxdot": v["x"].dot.data}


# -- C.6 and general-DAG cases -------------------------------------------------

def flux_1d(fd, n=11, n_t=6, seed=10):
    """Auxiliary edge field with n_time_ders=0 updated by assign_update."""
    m = fd.Model({"t": np.arange(n_t) * 0.1, "x": np.arange(n) * 1.0}, "t")
    T = fd.Field(m, "T", n_time_ders=1)
    q = fd.Field(m, "q", edge_axes=("x",))
    ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
    ec = fd.Stencil([-.5, .5], 1, "edge", "cell")
    _inject_ic(T, 0, _rng(seed).random(n))
    q.set_BC("0", "x", "start")
    q.set_BC("0", "x", "end")

    def body():
        q.assign_update(-1.0 * ce.der(T.prev, "x"))
        T.dot.assign_update(-1.0 * ec.der(q.new, "x"))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "q": q.data, "Tdot": T.dot.data}


def variable_coefficient(fd, nx=10, ny=7, n_t=5, seed=11):
    """ConstantField coefficient, set from an expression."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(nx) * 1.0,
                  "y": np.arange(ny) * 1.0}, "t", periodic=("x", "y"))
    T = fd.Field(m, "T", n_time_ders=1)
    kappa = fd.ConstantField(m, "kappa")
    kappa.set("1+x*y/50")
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(T, 0, _rng(seed).random((nx, ny)))

    def body():
        lap = d2.der(T.prev, "x") + d2.der(T.prev, "y")
        T.dot.assign_update(kappa.always * lap)
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data, "kappa": kappa.data}


def burgers_1d(fd, n=32, n_t=6, seed=12):
    """Nonlinear DAG: products of fields, negation, pow, division, rsub quirk."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(n) * 0.5}, "t",
                 periodic=("x",))
    u = fd.Field(m, "u", n_time_ders=1)
    d1 = fd.Stencil([-1, 0, 1], 1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    _inject_ic(u, 0, 1.0 + _rng(seed).random(n))

    def body():
        adv = -(u.prev * d1.der(u.prev, "x"))
        visc = 0.05 * d2.der(u.prev, "x")
        extra = (u.prev ** 2) / 100 + (1 - u.prev) * 1e-3 - 2.0 + u.prev ** 1.5 * 1e-4
        u.dot.assign_update(adv + visc + extra)
        u.time_integrate_update()

    _loop(m, body)
    return {"u": u.data, "udot": u.dot.data}


def edge_to_edge(fd, n=14, n_t=4, seed=13):
    """edge->edge stencil on a fixed edge field plus asymmetric points."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(n) * 1.0}, "t")
    f = fd.Field(m, "f", edge_axes="x", n_time_ders=1)
    dee = fd.Stencil([-1, 0, 1], 2, "edge", "edge")
    up = fd.Stencil([-2, -1, 0], 1, "edge", "edge")
    _inject_ic(f, 0, _rng(seed).random(n + 1))
    f.set_BC("0", "x", "start")
    f.set_BC("t", "x", "end")

    def body():
        f.dot.assign_update(0.1 * dee.der(f.prev, "x"))
        f.time_integrate_update()

    _loop(m, body)
    probe = up.der(f.prev, "x").data
    return {"f": f.data, "fdot": f.dot.data, "probe": np.array(probe)}


def cross_diffusion(fd, shape=(16, 6, 8), n_t=7, seed=17):
    """Two fields, and the second one's statement differentiates the FIRST
    field's current level as well: on several GPUs the ghost planes of `u` are
    read by two statements of the same step (slabs thick enough for the
    overlapped exchange at world sizes 2 and 4)."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(shape[0]) * 1.0, "y": np.arange(shape[1]) * 1.0,
                  "z": np.arange(shape[2]) * 1.0}, "t", periodic=("x", "y", "z"))
    u = fd.Field(m, "u", n_time_ders=1)
    c = fd.Field(m, "c", n_time_ders=1)
    d1 = fd.Stencil([-1, 0, 1], 1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    rng = _rng(seed)
    _inject_ic(u, 0, rng.random(shape))
    _inject_ic(c, 0, rng.random(shape))

    def body():
        u.dot.assign_update(0.1 * (d2.der(u.prev, "x") + d2.der(u.prev, "y") + d2.der(u.prev, "z")))
        c.dot.assign_update(0.05 * d1.der(u.prev, "x") + 0.02 * d2.der(c.prev, "x") + 0.02 * d2.der(c.prev, "z"))
        u.time_integrate_update()
        c.time_integrate_update()

    _loop(m, body)
    return {"u": u.data, "c": c.data, "cdot": c.dot.data}


def nested_divgrad(fd, n_t=5):
    """der of a computed expression, twice: div(kappa grad T) written with
    first-derivative stencils (FD.py:990-1003 accepts any FieldInstant)."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(12) * 1.0, "y": np.arange(16) * 0.5}, "t",
                 periodic=("x", "y"))
    T = fd.Field(m, "T", n_time_ders=1)
    kappa = fd.ConstantField(m, "kappa")
    kappa.set("1+x*y/50")
    d1 = fd.Stencil([-1, 0, 1], 1)
    T.set_IC("x*x+y")

    def body():
        T.dot.assign_update(d1.der(kappa.always * d1.der(T.prev, "x"), "x")
                            + d1.der(kappa.always * d1.der(T.prev, "y"), "y"))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def nested_burgers(fd, n_t=5):
    """Conservative Burgers flux: der of a product of stored levels, fixed axis
    with boundary values (the inner array has no NaN, the der of it does)."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(40) * 0.5}, "t")
    u = fd.Field(m, "u", n_time_ders=1)
    d1 = fd.Stencil([-1, 0, 1], 1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    u.set_IC("1+x/40")
    u.set_BC("1", "x", "start")
    u.set_BC("1.5", "x", "end")

    def body():
        u.dot.assign_update(-0.5 * d1.der(u.prev * u.prev, "x") + 0.05 * d2.der(u.prev, "x"))
        u.time_integrate_update()

    _loop(m, body)
    return {"u": u.data, "udot": u.dot.data}


def nested_staggered_flux(fd, n_t=5):
    """edge->cell der of (a factor times a cell->edge der): the flux form of
    diffusion without an auxiliary field; the inner array lives on the edges of a
    fixed axis (one more point than the field)."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(20) * 1.0, "y": np.arange(8) * 1.0}, "t",
                 periodic=("y",))
    T = fd.Field(m, "T", n_time_ders=1)
    ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
    ec = fd.Stencil([-.5, .5], 1, "edge", "cell")
    T.set_IC("x+y*y")
    T.set_BC("0", "x", "start")
    T.set_BC("3", "x", "end")

    def body():
        T.dot.assign_update(ec.der(0.3 * ce.der(T.prev, "x"), "x"))
        T.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


def nested_cross_axis(fd, n_t=5):
    """Ders of computed expressions that stay clear of the first axis on the
    outside (so that they can also be fused on slabs, where the first axis is
    the decomposed one): div_y(kappa grad_y T) + a plain x term, and a mixed
    second derivative d/dy(0.5 dc/dx) whose INNER der runs along x."""
    m = fd.Model({"t": np.arange(n_t) * 1e-2, "x": np.arange(12) * 1.0, "y": np.arange(16) * 0.5}, "t",
                 periodic=("x", "y"))
    T = fd.Field(m, "T", n_time_ders=1)
    c = fd.Field(m, "c", n_time_ders=1)
    kappa = fd.ConstantField(m, "kappa")
    kappa.set("1+x*y/50")
    d1 = fd.Stencil([-1, 0, 1], 1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    T.set_IC("x*x+y")
    c.set_IC("x*y+y*y")

    def body():
        T.dot.assign_update(d1.der(kappa.always * d1.der(T.prev, "y"), "y") + 0.1 * d2.der(T.prev, "x"))
        c.dot.assign_update(d1.der(0.5 * d1.der(c.prev, "x"), "y"))
        T.time_integrate_update()
        c.time_integrate_update()

    _loop(m, body)
    return {"T": T.data, "c": c.data, "cdot": c.dot.data}


def bc_conflict_prints(fd):
    """set_BC on top of set_IC prints the conflict note (FD.py:569-570)."""
    m = fd.Model({"t": range(0, 4), "x": range(0, 6)}, "t")
    T = fd.Field(m, "T", n_time_ders=1)
    T.set_IC("x*x")
    T.set_BC("3", "x", "start")
    T.set_BC("t", "x", "end")
    d = fd.Stencil([-1, 0, 1], 2)

    def body():
        T.dot.assign_update(0.1 * d.der(T.prev, "x"))
        T.time_integrate_update()

    m.check_IC()
    _loop(m, body)
    return {"T": T.data, "Tdot": T.dot.data}


ALL = {
    "readme_1d": readme_1d,
    "heat2d_fixed": heat2d_fixed,
    "heat2d_array_bc": heat2d_array_bc,
    "heat3d_periodic": heat3d_periodic,
    "heat3d_fixed": heat3d_fixed,
    "heat1d_odd_dx": heat1d_odd_dx,
    "heat4d_periodic": heat4d_periodic,
    "heat2d_time_middle": heat2d_time_middle,
    "wave2d": wave2d,
    "wave2d_5pt": wave2d_5pt,
    "wave2d_fixed_nan_spread": wave2d_fixed_nan_spread,
    "acoustic3d": acoustic3d,
    "acoustic3d_fixed": acoustic3d_fixed,
    "acoustic3d_wide": acoustic3d_wide,
    "acoustic2d": acoustic2d,
    "flux_1d": flux_1d,
    "variable_coefficient": variable_coefficient,
    "burgers_1d": burgers_1d,
    "edge_to_edge": edge_to_edge,
    "bc_conflict_prints": bc_conflict_prints,
    "cross_diffusion": cross_diffusion,
    "nested_divgrad": nested_divgrad,
    "nested_burgers": nested_burgers,
    "nested_staggered_flux": nested_staggered_flux,
    "nested_cross_axis": nested_cross_axis,
}
