"""CPU oracle for the finite_difference time-stepping hot path.

TEST INFRASTRUCTURE ONLY.  This module is a numpy restatement of the
reference algorithm (Ephraim-Bryski/Finite-Difference 0.0.3, file
``src/finite_difference/FD.py``; cited below as ``FD.py:<lines>``).  It is
imported only by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.  The shipped
package (``finite-difference_b200/finite_difference``) never imports it and has
no CPU fallback.

Parity pinning: the reference ships no tests, golden vectors or fixtures
(SURVEY.md section 4).  The oracle is therefore pinned against outputs of the
reference itself: ``oracle/make_golden.py`` imports the unmodified reference
from ``/root/reference`` (build container only), runs the scenarios of
SURVEY.md Appendix C and stores the results in ``tests/golden/*.npz``;
``tests/test_oracle.py`` checks this restatement bit-for-bit against those
fixtures everywhere and against the live reference where it is present.

The functions deliberately issue the same sequence of numpy array operations
as the reference (one ``np.roll`` + multiply + accumulate per stencil point,
boolean-mask merge, ...), so that timing them is a fair stand-in ("port") for
the reference's CPU path on machines where ``/root/reference`` is absent.
"""

from __future__ import annotations

import math
from dataclasses import dataclass, field as _dc_field

import numpy as np

MATH_NAMES = ("exp", "abs", "sin", "cos", "tan", "arcsin", "arccos", "arctan",
              "sinh", "cosh", "tanh")  # FD.py:405


# --------------------------------------------------------------------------
# stencil weights and shifts
# --------------------------------------------------------------------------

def stencil_weights(points, der_order):
    """Weights of the finite-difference approximation, rounded to 2 decimals.

    FD.py:938-951: Taylor matrix on doubled (half-step) coordinates,
    ``M[i, j] = (2 p_j) ** i``, right-hand side ``b[d] = d! * 2**d``,
    ``w = around(inv(M) b, 2)``.
    """
    doubled = np.array(points) * 2
    n = len(doubled)
    rows = np.arange(n).reshape(n, 1)
    taylor = np.power(np.asmatrix(doubled), np.asmatrix(rows))
    rhs = np.zeros((n, 1))
    rhs[der_order] = math.factorial(der_order) * 2 ** der_order
    solved = np.linalg.inv(taylor) * rhs
    return np.around(solved, 2).transpose().tolist()[0]


def stencil_kind(points):
    """True for half-valued points, False for integers (FD.py:909-920)."""
    pts = np.array(points)
    if np.all(pts % 1 == 0):
        return False
    if np.all((pts + 0.5) % 1 == 0):
        return True
    raise Exception("points need to be all integers or all half values")


def stencil_shifts(points, from_edge, to_edge):
    """Integer index shifts for one ``der`` call (FD.py:1028-1044).

    Returns ``(shifts, shape_shift)``.  Edge->cell moves the points by +1/2 and
    shortens a fixed axis by one; cell->edge moves them by -1/2 and lengthens
    it by one.  The float->int conversion truncates toward zero.
    """
    pts = np.array(points)
    if from_edge and not to_edge:
        return (pts + 1 / 2).astype(int), -1
    if (not from_edge) and to_edge:
        return (pts - 1 / 2).astype(int), 1
    return pts.astype(int), 0


# --------------------------------------------------------------------------
# der
# --------------------------------------------------------------------------

class Instability(Exception):
    pass


INSTABILITY_MSG = "Instability has led to infinite values, considering decreasing timestep"
OVERRIDE_MSG = "attempting to override values, this may be due to too many boundary conditions."


def der(f, points, weights, der_order, axis, periodic, from_edge, to_edge, dx):
    """Derivative of one time slice along ``axis`` (index among spatial axes).

    Periodic branch FD.py:1046-1055, fixed branch FD.py:1058-1105, scaling
    FD.py:1107-1110, infinity check FD.py:1112-1113.
    """
    shifts, shape_shift = stencil_shifts(points, from_edge, to_edge)
    if periodic:
        out = np.zeros(f.shape, dtype=f.dtype)   # f.dtype: float64 in the reference; float32 restates the fp32 option
        for w, s in zip(weights, shifts):
            out += w * np.roll(f, -s, axis=axis)
    else:
        out_shape = list(f.shape)
        out_shape[axis] += shape_shift
        out = np.full(tuple(out_shape), np.nan, dtype=f.dtype)
        right_guard = 1 if (from_edge and not to_edge) else 0
        lo_shift = min(shifts[0], 0)
        hi_shift = max(shifts[-1], right_guard)
        span = f.shape[axis] + lo_shift - hi_shift
        part_shape = list(out_shape)
        part_shape[axis] = span
        part = np.zeros(part_shape, dtype=f.dtype)
        sel = [slice(None)] * f.ndim
        for w, s in zip(weights, shifts):
            sel[axis] = slice(s - lo_shift, s - lo_shift + span)
            part += w * f[tuple(sel)]
        sel[axis] = slice(-lo_shift, -lo_shift + span)
        out[tuple(sel)] = part
    out /= f.dtype.type(dx ** der_order)
    if np.any(np.isinf(out)):
        raise Instability(INSTABILITY_MSG)
    return out


def der_scalar_1d(f, points, weights, der_order, periodic, from_edge, to_edge, dx):
    """Pure-Python scalar restatement of ``der`` for a 1-D slice (SURVEY.md
    Appendix A.2).  Slow; used as an independent cross-check on tiny inputs."""
    shifts, shape_shift = stencil_shifts(points, from_edge, to_edge)
    shifts = [int(s) for s in shifts]
    n = len(f)
    scale = dx ** der_order
    if periodic:
        out = [0.0] * n
        for j in range(n):
            acc = 0.0
            for w, s in zip(weights, shifts):
                acc = acc + w * float(f[(j + s) % n])
            out[j] = acc / scale
        return np.array(out, dtype=np.float64)
    n_out = n + shape_shift
    right_guard = 1 if (from_edge and not to_edge) else 0
    lo = -min(shifts[0], 0)
    hi = n - max(shifts[-1], right_guard)
    out = [float("nan")] * n_out
    for j in range(lo, hi):
        acc = 0.0
        for w, s in zip(weights, shifts):
            acc = acc + w * float(f[j + s])
        out[j] = acc / scale
    return np.array(out, dtype=np.float64)


# --------------------------------------------------------------------------
# merge-store and Euler update
# --------------------------------------------------------------------------

def merge(existing, new):
    """NaN-merge of freshly computed values into what is already stored.

    FD.py:556-576.  Returns ``(merged, overriding)``; ``new`` is modified in
    place exactly like the reference modifies its argument (FD.py:574).
    """
    unknown = np.isnan(new)
    known = np.invert(unknown)
    overriding = not np.all(np.isnan(existing[known]))
    new[unknown] = existing[unknown]
    return new, overriding


def euler(prev, dot_new, dt):
    """Forward-Euler update ``prev + dot_new*dt`` (FD.py:515-516)."""
    return prev + dot_new * dt


# --------------------------------------------------------------------------
# coordinates and IC/BC expressions
# --------------------------------------------------------------------------

def axis_coordinates(values, is_edge, is_periodic):
    """Sample coordinates of one axis (FD.py:415-441).  Edge ``i`` sits at
    ``x_i - dx/2``; a fixed edge axis has one extra point at the far end."""
    values = np.asarray(values)
    if not is_edge:
        return values
    dx = values[1] - values[0]
    coords = [v - dx / 2 for v in values]
    if not is_periodic:
        coords.append(coords[-1] + dx)
    return np.array(coords)


def evaluate_expression(expression, names, coords):
    """Evaluate an IC/BC string on the tensor product of ``coords``.

    Restates FD.py:446-471 with broadcasting instead of itertools.product +
    numexpr.  The identifier whitelist is the reference's (FD.py:402-408).
    numexpr itself is a third-party dependency (pyproject.toml:27, unpinned)
    that is not vendored in the reference: PARITY UNPINNED for transcendental
    functions at this single call site (FD.py:462); polynomial expressions are
    exact.
    """
    import re
    idents = [v for v in re.findall(r"\b[A-Za-z_][A-Za-z0-9_]*\b", expression)
              if v not in MATH_NAMES]
    unknown = set(idents) - set(names)
    if unknown:
        raise Exception(f"variables {unknown} are not axes names")
    env = {n: getattr(np, n) for n in MATH_NAMES}
    env["abs"] = np.abs
    env["__builtins__"] = {}
    local = {}
    nd = len(names)
    for i, (n, c) in enumerate(zip(names, coords)):
        shape = [1] * nd
        shape[i] = len(c)
        local[n] = np.asarray(c).reshape(shape)
    try:
        val = eval(expression, env, local)
    except Exception:
        raise Exception("cannot parse expression")
    full = tuple(len(c) for c in coords)
    return np.array(np.broadcast_to(np.asarray(val, dtype=np.float64), full))


# --------------------------------------------------------------------------
# a minimal run harness: full-history arrays like the reference (FD.py:310)
# --------------------------------------------------------------------------

@dataclass
class Grid:
    """Axes in user order; ``time`` names the time axis; time must be first
    for the harness (the benchmark layout)."""
    axes: dict
    time: str
    periodic: tuple = ()

    def __post_init__(self):
        self.axes = {k: np.array(v) for k, v in self.axes.items()}
        names = list(self.axes)
        assert names[0] == self.time, "oracle harness wants the time axis first"
        self.space = names[1:]
        self.step = 0

    @property
    def n_t(self):
        return len(self.axes[self.time])

    @property
    def dt(self):
        t = self.axes[self.time]
        return t[1] - t[0]

    def dx(self, axis):
        a = self.axes[axis]
        return a[1] - a[0]

    def shape(self, edge_axes=()):
        out = []
        for a in self.space:
            n = len(self.axes[a])
            if a in edge_axes and a not in self.periodic:
                n += 1
            out.append(n)
        return tuple(out)


@dataclass
class Series:
    """History of one field: ``data[t, ...]``, NaN = unknown (FD.py:310)."""
    grid: Grid
    edge_axes: tuple = ()
    data: np.ndarray = _dc_field(default=None)
    dtype: type = np.float64

    def __post_init__(self):
        if self.data is None:
            self.data = np.full((self.grid.n_t,) + self.grid.shape(self.edge_axes), np.nan, dtype=self.dtype)

    def set_face(self, axis, side, values):
        """Write a boundary face for all times (FD.py:361-385, 541-576 with
        allow_override=True: known new values replace, NaN new values keep)."""
        k = 1 + self.grid.space.index(axis)
        sel = [slice(None)] * self.data.ndim
        sel[k] = 0 if side == "start" else -1
        sel = tuple(sel)
        new = np.array(np.broadcast_to(values, self.data[sel].shape), dtype=self.data.dtype)
        merged, _ = merge(self.data[sel], new)
        self.data[sel] = merged

    def store(self, t, values):
        """Merge-store into time index ``t`` (FD.py:541-576, no override)."""
        merged, overriding = merge(self.data[t], values)
        if overriding:
            raise Exception(OVERRIDE_MSG)
        self.data[t] = merged


@dataclass
class StencilSpec:
    points: list
    order: int
    from_edge: bool = False
    to_edge: bool = False
    weights: list = None

    def __post_init__(self):
        if self.weights is None:
            self.weights = stencil_weights(self.points, self.order)

    def apply(self, grid: Grid, f, axis):
        return der(f, self.points, self.weights, self.order,
                   grid.space.index(axis), axis in grid.periodic,
                   self.from_edge, self.to_edge, grid.dx(axis))


def integrate(series: Series, dot: Series, grid: Grid):
    """time_integrate_update (FD.py:500-538) incl. the two NaN counts."""
    t = grid.step
    new = euler(series.data[t], dot.data[t + 1], series.data.dtype.type(grid.dt))
    series.store(t + 1, new)
    lost = int(np.isnan(series.data[t + 1]).sum()) > int(np.isnan(series.data[t]).sum())
    return lost and (t + 1) > 1


# ---- the five BASELINE.json workloads -------------

def run_heat(shape, n_t, periodic, k, dt, ic, bc_value=0.0, stencil_points=(-1, 0, 1),
             dx=1.0, keep_dot=True, dtype=np.float64):
    """C.1-C.3: dT/dt = k * sum_a d2T/da2.  ``shape`` is the spatial shape;
    fixed axes get Dirichlet ``bc_value`` on both faces.  Returns (T, Tdot)."""
    names = ["x", "y", "z", "w"][:len(shape)]
    axes = {"t": np.arange(n_t) * dt}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * dx
    g = Grid(axes, "t", tuple(a for a in names if periodic))
    T, Td = Series(g, dtype=dtype), Series(g, dtype=dtype)
    T.data[0] = ic
    if not periodic:
        for a in names:
            T.set_face(a, "start", bc_value)
            T.set_face(a, "end", bc_value)
    st = StencilSpec(list(stencil_points), 2)
    while g.step + 1 < g.n_t:
        prev = T.data[g.step]
        acc = st.apply(g, prev, names[0])
        for a in names[1:]:
            acc = acc + st.apply(g, prev, a)
        Td.store(g.step + 1, k * acc)
        integrate(T, Td, g)
        g.step += 1
    return T.data, Td.data


def run_wave(shape, n_t, c2, dt, ic_u, ic_v, stencil_points=tuple(range(-4, 5)), dx=1.0,
             periodic=True):
    """C.4: u'' = c2 * sum_a d2u/da2 with n_time_ders=2 (symplectic-Euler-like
    chain).  Returns (u, udot, udotdot)."""
    names = ["x", "y", "z"][:len(shape)]
    axes = {"t": np.arange(n_t) * dt}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * dx
    g = Grid(axes, "t", tuple(names) if periodic else ())
    U, Ud, Udd = Series(g), Series(g), Series(g)
    U.data[0] = ic_u
    Ud.data[0] = ic_v
    st = StencilSpec(list(stencil_points), 2)
    while g.step + 1 < g.n_t:
        prev = U.data[g.step]
        acc = st.apply(g, prev, names[0])
        for a in names[1:]:
            acc = acc + st.apply(g, prev, a)
        Udd.store(g.step + 1, c2 * acc)
        integrate(Ud, Udd, g)
        integrate(U, Ud, g)
        g.step += 1
    return U.data, Ud.data, Udd.data


def run_acoustic(shape, n_t, rho, K, dt, ic_p, dx=1.0, periodic=True, dtype=np.float64):
    """C.5: staggered pressure/velocity system.  p on cells, v_a on the edges
    of axis a.  Fixed variant: wall BC v_a = 0 on both end edges.
    Returns (p, [v_a...])."""
    names = ["x", "y", "z"][:len(shape)]
    axes = {"t": np.arange(n_t) * dt}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * dx
    g = Grid(axes, "t", tuple(names) if periodic else ())
    P, Pd = Series(g, dtype=dtype), Series(g, dtype=dtype)
    V = [Series(g, (a,), dtype=dtype) for a in names]
    Vd = [Series(g, (a,), dtype=dtype) for a in names]
    P.data[0] = ic_p
    for v, a in zip(V, names):
        v.data[0] = 0.0
        if not periodic:
            v.set_face(a, "start", 0.0)
            v.set_face(a, "end", 0.0)
    ce = StencilSpec([-0.5, 0.5], 1, False, True)
    ec = StencilSpec([-0.5, 0.5], 1, True, False)
    while g.step + 1 < g.n_t:
        for v, vd, a in zip(V, Vd, names):
            vd.store(g.step + 1, (-1 / rho) * ce.apply(g, P.data[g.step], a))
            integrate(v, vd, g)
        acc = ec.apply(g, V[0].data[g.step + 1], names[0])
        for v, a in zip(V[1:], names[1:]):
            acc = acc + ec.apply(g, v.data[g.step + 1], a)
        Pd.store(g.step + 1, (-K) * acc)
        integrate(P, Pd, g)
        g.step += 1
    return P.data, [v.data for v in V]
