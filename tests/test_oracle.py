"""The numpy oracle against the reference's own outputs (golden fixtures made
by oracle/make_golden.py from the unmodified reference) and, where
/root/reference is present, against the live reference."""

import json
import os

import numpy as np
import pytest

from oracle import fd_numpy as O
from oracle import ref_loader

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _same(a, b):
    return a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


@pytest.fixture(scope="module")
def scen():
    return np.load(os.path.join(GOLD, "scenarios.npz"))


def test_weights_table():
    table = json.load(open(os.path.join(GOLD, "weights.json")))
    assert len(table) >= 15
    for row in table:
        assert O.stencil_weights(row["points"], row["order"]) == row["weights"], row


def test_der_cases_bit_exact():
    cases = json.load(open(os.path.join(GOLD, "der_cases.json")))
    pack = np.load(os.path.join(GOLD, "der_cases.npz"))
    assert len(cases) == 120
    for i, c in enumerate(cases):
        fe, te = c["kind"][0] == "e", c["kind"][1] == "e"
        w = O.stencil_weights(c["points"], c["order"])
        assert w == c["weights"]
        out = O.der(pack[f"in{i}"], c["points"], w, c["order"], c["axis"], c["periodic"], fe, te, c["dx"])
        assert _same(out, pack[f"out{i}"]), c


def test_der_scalar_restatement_1d():
    cases = json.load(open(os.path.join(GOLD, "der_cases.json")))
    pack = np.load(os.path.join(GOLD, "der_cases.npz"))
    n = 0
    for i, c in enumerate(cases):
        if c["ndim"] != 1:
            continue
        fe, te = c["kind"][0] == "e", c["kind"][1] == "e"
        out = O.der_scalar_1d(pack[f"in{i}"], c["points"], c["weights"], c["order"], c["periodic"], fe, te, c["dx"])
        assert _same(out, pack[f"out{i}"]), c
        n += 1
    assert n >= 30


def test_heat2d_fixed(scen):
    ic = np.random.default_rng(0).random((12, 10))
    T, Td = O.run_heat((12, 10), 6, False, 0.1, 1e-3, ic)
    assert _same(T, scen["heat2d_fixed/T"]) and _same(Td, scen["heat2d_fixed/Tdot"])


def test_heat3d_periodic(scen):
    ic = np.random.default_rng(0).random((6, 5, 4))
    T, Td = O.run_heat((6, 5, 4), 7, True, 0.1, 1e-3, ic)
    assert _same(T, scen["heat3d_periodic/T"]) and _same(Td, scen["heat3d_periodic/Tdot"])
    # mean conserved
    assert abs(T[-1].mean() - T[0].mean()) < 1e-15


def test_wave2d(scen):
    ic = np.random.default_rng(5).random((16, 12))
    u, ud, udd = O.run_wave((16, 12), 6, 0.01, 1e-2, ic, np.zeros((16, 12)))
    assert _same(u, scen["wave2d/u"]) and _same(ud, scen["wave2d/udot"]) and _same(udd, scen["wave2d/udotdot"])


def test_wave2d_fixed_nan_spread(scen):
    ic = np.random.default_rng(7).random((24, 24))
    u, ud, udd = O.run_wave((24, 24), 4, 0.01, 1e-2, ic, np.zeros((24, 24)),
                            stencil_points=(-2, -1, 0, 1, 2), periodic=False)
    assert _same(u, scen["wave2d_fixed_nan_spread/u"])
    # NaN front advances by the radius (2) per step
    assert np.isnan(u[1]).sum() < np.isnan(u[2]).sum() < np.isnan(u[3]).sum()


@pytest.mark.parametrize("periodic,name,shape,seed", [(True, "acoustic3d", (6, 5, 4), 8),
                                                      (False, "acoustic3d_fixed", (5, 6, 4), 9)])
def test_acoustic(scen, periodic, name, shape, seed):
    ic = np.random.default_rng(seed).random(shape)
    p, vs = O.run_acoustic(shape, 6, 1.0, 0.25, 0.1, ic, periodic=periodic)
    assert _same(p, scen[f"{name}/p"])
    for v, a in zip(vs, "xyz"):
        assert _same(v, scen[f"{name}/v{a}"])
    assert abs(p[-1].sum() - p[0].sum()) < 1e-12


def test_axis_coordinates_and_expression():
    c = O.axis_coordinates(np.arange(5) * 2.0, True, False)
    assert np.array_equal(c, [-1, 1, 3, 5, 7, 9])
    c = O.axis_coordinates(np.arange(5) * 2.0, True, True)
    assert np.array_equal(c, [-1, 1, 3, 5, 7])
    v = O.evaluate_expression("x*y+1", ["x", "y"], [np.arange(3.0), np.arange(2.0)])
    assert np.array_equal(v, [[1, 1], [1, 2], [1, 3]])
    with pytest.raises(Exception, match="not axes names"):
        O.evaluate_expression("sqrt(x)", ["x"], [np.arange(3.0)])


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present on this machine")
def test_oracle_vs_live_reference_random_der():
    """400 extra random 1-D/2-D cases against the reference imported live."""
    import contextlib
    import io
    fd = ref_loader.load()
    rng = np.random.default_rng(99)
    for i in range(400):
        kind = ["cc", "ce", "ec", "ee"][i % 4]
        half = kind in ("ce", "ec")
        npts = int(rng.integers(2, 7))
        start = int(rng.integers(-4, 2))
        pts = [start + j + (0.5 if half else 0) for j in range(npts)]
        if not half:
            pts = [int(p) for p in pts]
        order = int(rng.integers(1, npts))
        periodic = bool(rng.integers(0, 2))
        n, n2 = int(rng.integers(8, 14)), int(rng.integers(2, 5))
        dx = float(rng.choice([1.0, 0.25, 0.3, 3.0]))
        fe, te = kind[0] == "e", kind[1] == "e"
        m = fd.Model({"t": [0, 1], "y": np.arange(n2) * 1.0, "x": np.arange(n) * dx}, "t",
                     periodic=("x",) if periodic else ())
        f = fd.Field(m, "f", edge_axes=("x",) if fe else (), n_time_ders=1)
        with contextlib.redirect_stdout(io.StringIO()):
            st = fd.Stencil(pts, order, "edge" if fe else "cell", "edge" if te else "cell")
        vals = rng.standard_normal(f.data.shape[1:])
        f.data[0] = vals
        ref = st.der(f.prev, "x").data
        mine = O.der(vals, pts, O.stencil_weights(pts, order), order, 1, periodic, fe, te, m.axes["x"][1] - m.axes["x"][0])
        assert _same(np.asarray(ref), mine), (pts, order, kind, periodic)


@pytest.mark.skipif(not ref_loader.available(), reason="reference source not present on this machine")
def test_port_runs_at_the_speed_of_the_live_reference():
    """bench.py falls back to timing the numpy port where the reference file is
    absent; this measures that the port is a fair stand-in: the same 128^3 x 3-step
    heat run through the unmodified FD.py and through the port, interleaved, best of
    five each (measured: 0.9-1.1; the bound of 25 % leaves room for a loaded
    host), and the same bits."""
    import bench
    shape = (128, 128, 128)
    ic = np.random.default_rng(5).random(shape)
    kind, live = bench.reference_step_fn("c3", shape)
    assert kind == "reference"
    saved = ref_loader.available
    ref_loader.available = lambda: False
    try:
        kind, port = bench.reference_step_fn("c3", shape)
    finally:
        ref_loader.available = saved
    assert kind == "port"
    t_live, t_port = [], []
    for _ in range(5):
        s, state_live = live(3, (ic, None))
        t_live.append(s)
        s, state_port = port(3, (ic, None))
        t_port.append(s)
    assert np.array_equal(state_live[0], state_port[0])
    ratio = min(t_port) / min(t_live)
    assert 0.75 < ratio < 1.33, (t_port, t_live)
