"""Heat equation on grids with more than three spatial axes (the reference's
der works on any axis of an N-D array, FD.py:1007-1012), shared by the CPU
twin (numpy stand-in for the library) and the GPU test."""

import contextlib
import io

import numpy as np

from oracle import fd_numpy as O

CASES = [((3, 4, 2, 5, 6), True, 4), ((4, 5, 3, 6), True, 4), ((2, 3, 4, 8), True, 3), ((5, 4, 6, 7), False, 4),
         ((3, 2, 3, 2, 4, 16), True, 4)]


def run_case(fd, shape, periodic, n_ders, n_t=4):
    """Returns (T.data of the package, T.data of the oracle, engine stats)."""
    names = ["x", "y", "z", "w", "v", "u"][:len(shape)]
    axes = {"t": np.arange(n_t) * 1e-3}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    ic = np.random.default_rng(len(shape)).random(shape)
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(dict(axes), "t", periodic=tuple(names) if periodic else ())
        T = fd.Field(m, "T", n_time_ders=1)
        d2 = fd.Stencil([-1, 0, 1], 2)
        T.set_IC_values(ic)
        if not periodic:
            for a in names:
                T.set_BC("0", a, "start")
                T.set_BC("0", a, "end")
        while not m.finished:
            acc = d2.der(T.prev, names[0])
            for a in names[1:n_ders]:
                acc = acc + d2.der(T.prev, a)
            T.dot.assign_update(0.1 * acc)
            T.time_integrate_update()
            m.increment_time()
    g = O.Grid(dict(axes), "t", tuple(names) if periodic else ())
    S, Sd = O.Series(g), O.Series(g)
    S.data[0] = ic
    if not periodic:
        for a in names:
            S.set_face(a, "start", 0.0)
            S.set_face(a, "end", 0.0)
    st = O.StencilSpec([-1, 0, 1], 2)
    while g.step + 1 < g.n_t:
        prev = S.data[g.step]
        acc = st.apply(g, prev, names[0])
        for a in names[1:n_ders]:
            acc = acc + st.apply(g, prev, a)
        Sd.store(g.step + 1, 0.1 * acc)
        O.integrate(S, Sd, g)
        g.step += 1
    return T.data, S.data, dict(m._engine.stats)
