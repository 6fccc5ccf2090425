"""Parity tests proper (B200): the package, through the C ABI of
libfd_b200.so, against the reference's golden outputs and the numpy oracle.
fp64 results must be bit-identical (NaN pattern included)."""

import contextlib
import ctypes as C
import io
import json
import os

import numpy as np
import pytest

from oracle import fd_numpy as O
from oracle import scenarios
from tests.scenario_check import GOLD, run_and_compare

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fd():
    import finite_difference as fd
    return fd


@pytest.mark.parametrize("name", sorted(scenarios.ALL))
def test_scenario(fd, name):
    run_and_compare(name, fd)


@pytest.mark.parametrize("name", sorted(scenarios.ALL))
def test_scenario_unfused_kernels(fd, name, monkeypatch):
    monkeypatch.setenv("FD_B200_NO_FUSE", "1")
    run_and_compare(name, fd)


def test_der_cases_through_the_c_abi():
    """fd_der_f64 called directly with plain pointers on the 120 golden cases."""
    import torch
    import finite_difference._native as nat
    cases = json.load(open(os.path.join(GOLD, "der_cases.json")))
    pack = np.load(os.path.join(GOLD, "der_cases.npz"))
    flags = torch.zeros(nat.FD_NUM_FLAGS, dtype=torch.int64, device="cuda")
    for i, c in enumerate(cases):
        fe, te = c["kind"][0] == "e", c["kind"][1] == "e"
        shifts, grow = O.stencil_shifts(c["points"], fe, te)
        vin, want = pack[f"in{i}"], pack[f"out{i}"]
        ax = c["axis"]
        n_src = vin.shape[ax]
        outer = int(np.prod(vin.shape[:ax])) if ax else 1
        inner = int(np.prod(vin.shape[ax + 1:])) if ax + 1 < vin.ndim else 1
        if c["periodic"]:
            n_dst, lo, hi = n_src, 0, n_src
        else:
            n_dst = n_src + grow
            guard = 1 if (fe and not te) else 0
            lo = -min(int(shifts[0]), 0)
            hi = lo + n_src + min(int(shifts[0]), 0) - max(int(shifts[-1]), guard)
        st = nat.make_stencil(shifts, c["weights"], c["dx"] ** c["order"])
        src = torch.from_numpy(vin.reshape(-1).copy()).cuda()
        dst = torch.empty(outer * n_dst * inner, dtype=torch.float64, device="cuda")
        nat.check(nat.lib.fd_der_f64(dst.data_ptr(), src.data_ptr(), outer, n_src, n_dst, inner, C.byref(st),
                                     1 if c["periodic"] else 0, lo, hi, flags.data_ptr(), 0))
        got = dst.cpu().numpy().reshape(want.shape)
        assert np.array_equal(got, want, equal_nan=True), c
    assert int(flags[nat.FD_FLAG_INF]) == 0


def _heat(fd, shape, n_t, periodic, k, dt, ic, points=(-1, 0, 1), history="all"):
    names = ["x", "y", "z"][:len(shape)]
    axes = {"t": np.arange(n_t) * dt}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=tuple(names) if periodic else (), history=history)
        T = fd.Field(m, "T", n_time_ders=1)
        d2 = fd.Stencil(list(points), 2)
        T.set_IC_values(ic)
        if not periodic:
            for a in names:
                T.set_BC("0", a, "start")
                T.set_BC("0", a, "end")
        while not m.finished:
            acc = d2.der(T.prev, names[0])
            for a in names[1:]:
                acc = acc + d2.der(T.prev, a)
            T.dot.assign_update(k * acc)
            T.time_integrate_update()
            m.increment_time()
    return m, T


@pytest.mark.parametrize("shape,periodic", [((67, 45, 130), True), ((33, 129, 70), False), ((1000, 37), False),
                                            ((513, 300), True), ((5000,), True), ((3, 3, 3), True),
                                            ((2, 2), False), ((40, 130, 520), False), ((700, 1024), False),
                                            ((3000,), False)])
def test_heat_against_oracle_odd_shapes(fd, shape, periodic):
    """Shapes that do not divide the thread-block tiles, both boundary kinds."""
    ic = np.random.default_rng(11).random(shape)
    m, T = _heat(fd, shape, 4, periodic, 0.1, 1e-3, ic)
    want, want_dot = O.run_heat(shape, 4, periodic, 0.1, 1e-3, ic)
    assert np.array_equal(T.data, want, equal_nan=True)
    assert np.array_equal(T.dot.data, want_dot, equal_nan=True)
    assert m._engine.stats["fused"] == 3 and m._engine.stats["unfused_statements"] == 0


def test_wide_stencil_against_oracle(fd):
    shape = (40, 52, 36)
    ic = np.random.default_rng(12).random(shape)
    m, T = _heat(fd, shape, 3, True, 0.05, 1e-3, ic, points=tuple(range(-4, 5)))
    want, _ = O.run_heat(shape, 3, True, 0.05, 1e-3, ic, stencil_points=tuple(range(-4, 5)))
    assert np.array_equal(T.data, want, equal_nan=True)


def test_history_policies(fd):
    shape = (20, 24, 28)
    ic = np.random.default_rng(13).random(shape)
    want, _ = O.run_heat(shape, 8, True, 0.1, 1e-3, ic)
    _, T_last = _heat(fd, shape, 8, True, 0.1, 1e-3, ic, history="last")
    got = T_last.data
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[-1], want[-1])
    assert np.isnan(got[1:-1]).all()
    _, T_k = _heat(fd, shape, 8, True, 0.1, 1e-3, ic, history=3)
    got = T_k.data
    for t in range(8):
        if t % 3 == 0 or t == 7:
            assert np.array_equal(got[t], want[t])
        else:
            assert np.isnan(got[t]).all()


@pytest.mark.parametrize("plant", ["nan", "inf"])
def test_lean_kernel_two_stage_nan_and_inf(fd, monkeypatch, plant):
    """The lean star kernel looks for NaN / infinite values through a magnitude
    sum first and exactly only when that sum is not finite: the growth of a NaN
    region (values and the 'losing data' note) and the instability error are
    those of the operator-by-operator path."""
    shape, n_t = (24, 32, 64), 6
    ic = np.random.default_rng(14).random(shape)
    ic[5, 6, 7] = np.nan if plant == "nan" else np.inf
    outcomes = []
    for unfused in (False, True):
        if unfused:
            monkeypatch.setenv("FD_B200_NO_FUSE", "1")
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            try:
                axes = {"t": np.arange(n_t) * 1e-3}
                axes.update({a: np.arange(n) * 1.0 for a, n in zip("xyz", shape)})
                m = fd.Model(axes, "t", periodic=("x", "y", "z"))
                T = fd.Field(m, "T", n_time_ders=1)
                d2 = fd.Stencil([-1, 0, 1], 2)
                T.set_IC_values(ic)
                while not m.finished:
                    T.dot.assign_update(0.1 * ((d2.der(T.prev, "x") + d2.der(T.prev, "y")) + d2.der(T.prev, "z")))
                    T.time_integrate_update()
                    m.increment_time()
                data = T.data
                outcomes.append(("ok", buf.getvalue().count("losing data"), data))
                assert unfused or m._engine.stats["fused"] == n_t - 1
            except AssertionError:
                raise
            except Exception as e:  # noqa: BLE001 - the message is what is compared
                outcomes.append((type(e).__name__, str(e), None))
    assert outcomes[0][:2] == outcomes[1][:2]
    if plant == "nan":
        want, _ = O.run_heat(shape, n_t, True, 0.1, 1e-3, ic)
        assert outcomes[0][0] == "ok" and outcomes[0][1] >= 1
        assert np.array_equal(outcomes[0][2], want, equal_nan=True)
    else:
        assert "Instability has led to infinite values" in outcomes[0][1]


def test_instability_and_override_are_reported(fd):
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model({"t": np.arange(40) * 1.0, "x": np.arange(16) * 1e-3}, "t", periodic=("x",))
        T = fd.Field(m, "T", n_time_ders=1)
        d2 = fd.Stencil([-1, 0, 1], 2)
        T.set_IC_values(np.random.default_rng(0).random(16) * 1e300)
        with pytest.raises(Exception, match="Instability has led to infinite values"):
            while not m.finished:
                T.dot.assign_update(1e8 * d2.der(T.prev, "x"))
                T.time_integrate_update()
                m.increment_time()
        # too many boundary conditions: BC on the derivative field as well
        m = fd.Model({"t": np.arange(4) * 1.0, "x": np.arange(8) * 1.0}, "t")
        u = fd.Field(m, "u", n_time_ders=1)
        u.set_IC("x")
        u.set_BC("0", "x", "start")
        u.set_BC("0", "x", "end")
        d1 = fd.Stencil([0, 1], 1)
        with pytest.raises(Exception, match="attempting to override values"):
            while not m.finished:
                u.dot.assign_update(d1.der(u.prev, "x"))
                u.time_integrate_update()
                m.increment_time()


def test_general_fused_kernel_without_star_specialisation():
    """FD_B200_NO_STAR=1 (read once per process) forces the general fused
    kernel for statements the star kernel would otherwise take."""
    import subprocess
    import sys
    code = (
        "import sys, os; sys.path[:0] = [%r, %r]\n"
        "import finite_difference as fd\n"
        "from tests.scenario_check import run_and_compare\n"
        "for name in ['readme_1d','heat2d_fixed','heat3d_periodic','heat3d_fixed','wave2d','heat2d_time_middle','variable_coefficient']:\n"
        "    run_and_compare(name, fd)\n"
        "print('ok')\n"
    ) % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
         os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "finite-difference_b200"))
    for switch in ("FD_B200_NO_STAR", "FD_B200_NO_LEAN"):   # general fused kernel; general star kernel
        env = dict(os.environ, **{switch: "1"})
        out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert out.returncode == 0 and out.stdout.strip().endswith("ok"), switch + out.stdout + out.stderr


# ---------------------------------------------------------------- fp32 option
def _acoustic(fd, shape, n_t, periodic, ic, dtype):
    names = ("x", "y", "z")
    axes = {"t": np.arange(n_t) * 0.1}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=names if periodic else (), dtype=dtype)
        p = fd.Field(m, "p", n_time_ders=1)
        v = {a: fd.Field(m, "v" + a, edge_axes=(a,), n_time_ders=1) for a in names}
        d_ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
        d_ec = fd.Stencil([-.5, .5], 1, "edge", "cell")
        p.set_IC_values(ic)
        for a in names:
            v[a].set_IC("0")
            if not periodic:
                v[a].set_BC("0", a, "start")
                v[a].set_BC("0", a, "end")
        while not m.finished:
            for a in names:
                v[a].dot.assign_update((-1 / 1.0) * d_ce.der(p.prev, a))
                v[a].time_integrate_update()
            div = d_ec.der(v["x"].new, "x") + d_ec.der(v["y"].new, "y") + d_ec.der(v["z"].new, "z")
            p.dot.assign_update((-0.25) * div)
            p.time_integrate_update()
            m.increment_time()
    return m, p, v


@pytest.mark.parametrize("periodic", [True, False])
def test_fp32_staggered_acoustic(fd, periodic):
    """fp32 option (not in the reference): bit-identical to the float32 numpy
    restatement, and within the stated bound of the fp64 oracle: relative L2
    error <= 1e-5 and max-abs error <= 1e-4 * max|p|."""
    shape, n_t = (36, 40, 64), 21
    ic64 = np.random.default_rng(31).random(shape)
    ic32 = ic64.astype(np.float32)
    m, p, v = _acoustic(fd, shape, n_t, periodic, ic32, np.float32)
    got = p.data
    want32, _ = O.run_acoustic(shape, n_t, 1.0, 0.25, 0.1, ic32, periodic=periodic, dtype=np.float32)
    assert np.array_equal(got, want32.astype(np.float64), equal_nan=True)
    want64, _ = O.run_acoustic(shape, n_t, 1.0, 0.25, 0.1, ic32.astype(np.float64), periodic=periodic)
    err = got[-1] - want64[-1]
    assert np.linalg.norm(err) / np.linalg.norm(want64[-1]) <= 1e-5
    assert np.abs(err).max() <= 1e-4 * np.abs(want64[-1]).max()
    assert m._engine.stats["unfused_statements"] == 0
    if periodic:   # the whole step is one fused velocity+pressure launch
        assert m._engine.stats.get("fused_pairs") == n_t - 1 and m._engine.stats["fused"] == 0


def test_fp64_staggered_pair_kernel(fd):
    shape, n_t = (20, 24, 130), 9
    ic = np.random.default_rng(33).random(shape)
    m, p, v = _acoustic(fd, shape, n_t, True, ic, np.float64)
    want, want_v = O.run_acoustic(shape, n_t, 1.0, 0.25, 0.1, ic)
    assert np.array_equal(p.data, want, equal_nan=True)
    for a, w in zip("xyz", want_v):
        assert np.array_equal(v[a].data, w, equal_nan=True)
    assert m._engine.stats.get("fused_pairs") == n_t - 1


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("variant", ["default", "march_vy1", "march_vy2", "no_march"])
def test_staggered_marching_pair_kernel(fd, monkeypatch, dtype, variant):
    """The marching pair kernel (large periodic 3-D staggered models; csrc/fd_pair_march.cuh),
    the plain pair kernel and the oracle agree bit for bit; the grid is large
    enough for the library to pick the marching kernel on its own."""
    if variant == "no_march":
        monkeypatch.setenv("FD_B200_NO_MARCH", "1")
    elif variant != "default":
        monkeypatch.setenv("FD_B200_PAIR_CHUNK", "8")
        monkeypatch.setenv("FD_B200_PAIR_VY", variant[-1])
    shape, n_t = (64, 128, 512), 4
    ic = np.random.default_rng(35).random(shape).astype(dtype)
    m, p, v = _acoustic(fd, shape, n_t, True, ic, dtype)
    want, want_v = O.run_acoustic(shape, n_t, 1.0, 0.25, 0.1, ic, dtype=dtype)
    assert np.array_equal(p.data, want.astype(np.float64), equal_nan=True)
    for a, w in zip("xyz", want_v):
        assert np.array_equal(v[a].data, w.astype(np.float64), equal_nan=True)
    assert m._engine.stats.get("fused_pairs") == n_t - 1


def test_marching_pair_kernel_reports_nan_and_inf(fd, monkeypatch):
    """NaN counts and the infinite-der flag of the marching kernel come from its
    two-stage test (finite magnitude sum, else exact): same deferred messages
    as the plain pair kernel."""
    import contextlib, io
    outcomes = []
    for march in (True, False):
        if march:
            monkeypatch.setenv("FD_B200_PAIR_CHUNK", "8")
        else:
            monkeypatch.delenv("FD_B200_PAIR_CHUNK")
            monkeypatch.setenv("FD_B200_NO_MARCH", "1")
        shape, n_t = (12, 16, 64), 4
        ic = np.random.default_rng(36).random(shape)
        ic[3, 4, 5] = np.inf
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            try:
                m, p, v = _acoustic(fd, shape, n_t, True, ic, np.float64)
                _ = p.data
                outcomes.append(("ok", buf.getvalue()))
            except Exception as e:  # noqa: BLE001 - the message is what is compared
                outcomes.append((type(e).__name__, str(e)))
    assert outcomes[0] == outcomes[1]
    assert outcomes[0][0] != "ok" and "infinite" in outcomes[0][1]


def test_fp32_heat_star_kernel(fd):
    shape = (24, 40, 512)
    ic32 = np.random.default_rng(32).random(shape).astype(np.float32)
    names = ("x", "y", "z")
    axes = {"t": np.arange(6) * 1e-3}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=names, dtype=np.float32)
        T = fd.Field(m, "T", n_time_ders=1)
        d2 = fd.Stencil([-1, 0, 1], 2)
        T.set_IC_values(ic32)
        while not m.finished:
            T.dot.assign_update(0.1 * (d2.der(T.prev, "x") + d2.der(T.prev, "y") + d2.der(T.prev, "z")))
            T.time_integrate_update()
            m.increment_time()
    want, _ = O.run_heat(shape, 6, True, 0.1, 1e-3, ic32, dtype=np.float32)
    assert np.array_equal(T.data, want.astype(np.float64), equal_nan=True)


def test_history_all_streams_every_level_correctly(fd):
    """Default policy (every level of every field comes back, like the
    reference) on a grid large enough that the device-to-host copies of a level
    are still in flight when the ring slot is due to be overwritten."""
    shape, n_t = (160, 128, 192), 10
    ic = np.random.default_rng(41).random(shape)
    m, T = _heat(fd, shape, n_t, True, 0.1, 1e-3, ic, history="all")
    want, want_dot = O.run_heat(shape, n_t, True, 0.1, 1e-3, ic)
    assert np.array_equal(T.data, want, equal_nan=True)
    assert np.array_equal(T.dot.data, want_dot, equal_nan=True)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_invariant_division_is_correctly_rounded(dtype):
    """The FMA-corrected reciprocal division (exact_div=1, div_mode 3) against
    the IEEE division sequence (exact_div=2) on ~10^7 values per divisor,
    including zeros, infinities, NaN, subnormals and huge magnitudes."""
    import torch
    import finite_difference._native as nat
    rng = np.random.default_rng(77)
    n = 1 << 20
    mags = rng.uniform(-300 if dtype == np.float64 else -37, 300 if dtype == np.float64 else 37, n)
    x = (rng.standard_normal(n) * 10.0 ** mags).astype(dtype)
    x[:16] = [0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, -5e-324, 1e-310, 1.0, -1.0, 3.0, 1e308, -1e308, 2.0 ** -1022, 0.1, 7.0]
    more = rng.standard_normal(n).astype(dtype)
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    fn = nat.lib.fd_der_f64 if dtype == np.float64 else nat.lib.fd_der_f32
    flags = torch.zeros(nat.FD_NUM_FLAGS, dtype=torch.int64, device="cuda")
    for scale in [0.1, 0.01, 0.3, 3.0, 7.0, 1e-3, 1.0 / 3.0, 0.09, 0.010000000000000002, 1e-6, 12345.678, 0.7 ** 2,
                  -0.3, 1e8]:
        for vals in (x, more):
            src = torch.from_numpy(vals.copy()).cuda()
            outs = []
            for exact in (1, 2):
                st = nat.make_stencil([0], [1.0], scale)
                st.exact_div = exact
                dst = torch.empty(n, dtype=tdt, device="cuda")
                nat.check(fn(dst.data_ptr(), src.data_ptr(), 1, n, n, 1, C.byref(st), 1, 0, n, flags.data_ptr(), 0))
                outs.append(dst.cpu().numpy())
            same = (outs[0].view(np.uint64 if dtype == np.float64 else np.uint32) ==
                    outs[1].view(np.uint64 if dtype == np.float64 else np.uint32)) | (np.isnan(outs[0]) & np.isnan(outs[1]))
            assert same.all(), (scale, vals[~same][:5], outs[0][~same][:5], outs[1][~same][:5])
            with np.errstate(all="ignore"):
                want = vals / dtype(scale)
            assert np.array_equal(outs[0], want, equal_nan=True), scale


def test_local_level_into_caller_buffer(fd):
    shape = (16, 20, 24)
    ic = np.random.default_rng(51).random(shape)
    m, T = _heat(fd, shape, 5, True, 0.1, 1e-3, ic, history="last")
    want, _ = O.run_heat(shape, 5, True, 0.1, 1e-3, ic)
    out = np.empty(shape)
    got = T.local_level(4, out=out)
    assert got is out and np.array_equal(out, want[4])
    assert np.array_equal(T.local_level(4), want[4])


def _wave(fd, shape, n_t, periodic, ic, points, dx=1.0, dtype=np.float64):
    names = ["x", "y"]
    axes = {"t": np.arange(n_t) * 1e-2}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * dx
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=tuple(names) if periodic else (), dtype=dtype)
        u = fd.Field(m, "u", n_time_ders=2)
        st = fd.Stencil(list(points), 2)
        u.set_IC_values(ic)
        u.dot.set_IC("0")
        while not m.finished:
            u.dot.dot.assign_update(0.01 * (st.der(u.prev, "x") + st.der(u.prev, "y")))
            u.dot.time_integrate_update()
            u.time_integrate_update()
            m.increment_time()
    return m, u


@pytest.mark.parametrize("shape,points,dx", [((100, 200), range(-4, 5), 1.0), ((37, 1000), range(-4, 5), 1.0),
                                             ((130, 136), range(-3, 4), 1.0), ((64, 4096), range(-4, 5), 0.5),
                                             ((75, 520), range(-4, 5), 0.3), ((300, 64), range(-3, 4), 1.0)])
def test_wide_2d_wave_row_kernel(fd, shape, points, dx):
    """Radius-3/4 stencils on 2-D periodic grids take the warp-shuffle row kernel
    (fd_star_row.cuh): runs that do not divide into whole warps, wrapped runs,
    every scaling class, depth-2 Euler chain, .dot.dot kept (history='all')."""
    import finite_difference._native as nat
    ic = np.random.default_rng(41).random(shape)
    m, u = _wave(fd, shape, 5, True, ic, points, dx)
    assert nat.lib.fd_last_kernel().decode().startswith("star_row f64"), nat.lib.fd_last_kernel()
    want_u, want_v, want_a = O.run_wave(shape, 5, 0.01, 1e-2, ic, np.zeros(shape), stencil_points=tuple(points), dx=dx)
    assert np.array_equal(u.data, want_u, equal_nan=True)
    assert np.array_equal(u.dot.data, want_v, equal_nan=True)
    assert np.array_equal(u.dot.dot.data, want_a, equal_nan=True)
    assert m._engine.stats["unfused_statements"] == 0


def test_wide_2d_wave_row_kernel_matches_general_kernel_fp32(fd):
    """fp32 (4-element vectors, one feeding lane per side): same bits as the
    general star kernel, which the fp32 tests pin against the float32 oracle."""
    shape = (90, 640)
    ic = np.random.default_rng(42).random(shape).astype(np.float32)
    _, u = _wave(fd, shape, 4, True, ic, range(-4, 5), dtype=np.float32)
    got = u.data
    # (FD_B200_NO_ROW is read once per process by the library: the general kernel runs in a subprocess)
    import subprocess
    import sys
    code = ("import sys, numpy as np\nsys.path[:0] = %r\nimport tests.test_gpu_parity as t, finite_difference as fd\n"
            "ic = np.random.default_rng(42).random((90, 640)).astype(np.float32)\n"
            "_, u = t._wave(fd, (90, 640), 4, True, ic, range(-4, 5), dtype=np.float32)\n"
            "import finite_difference._native as nat\nassert nat.lib.fd_last_kernel().decode().startswith('star f32'), nat.lib.fd_last_kernel()\n"
            "np.save(sys.argv[1], u.data)\n") % (sys.path[:3],)
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        out = os.path.join(tmp, "u.npy")
        subprocess.check_call([sys.executable, "-c", code, out], env=dict(os.environ, FD_B200_NO_ROW="1"))
        want = np.load(out)
    assert np.array_equal(got, want, equal_nan=True)


def test_wide_2d_heat_fixed_interior_on_row_kernel(fd):
    """Fixed boundaries: the interior box (clean) runs the row kernel, the rim the
    bounded general kernel; depth 1 with a 7-point stencil."""
    shape = (150, 700)
    ic = np.random.default_rng(43).random(shape)
    m, T = _heat(fd, shape, 4, False, 0.05, 1e-3, ic, points=tuple(range(-3, 4)))
    want, want_dot = O.run_heat(shape, 4, False, 0.05, 1e-3, ic, stencil_points=tuple(range(-3, 4)))
    assert np.array_equal(T.data, want, equal_nan=True)
    assert np.array_equal(T.dot.data, want_dot, equal_nan=True)


def test_more_than_three_spatial_axes_fused(fd):
    """4-, 5- and 6-D grids: the leading axes merge into dimension 0 of the
    kernels' view and a der along one of them carries its fold; one launch per
    statement, the oracle's bits.  (Fixed boundaries on merged axes keep the
    unfused operators.)"""
    from tests.nd_cases import CASES, run_case
    for shape, periodic, n_ders in CASES:
        got, want, stats = run_case(fd, shape, periodic, n_ders)
        assert np.array_equal(got, want, equal_nan=True), (shape, periodic)
        if periodic:
            assert stats["fused"] == 3 and stats["unfused_statements"] == 0, (shape, stats)


def test_general_power_is_within_the_stated_bound():
    """`**` with a general exponent is CUDA's pow (not correctly rounded, as is
    numpy's libm pow): not bit-identical to the reference, but inside the 1e-12
    relative bound of the north star -- in practice a couple of ulp.  numpy's
    scalar-exponent fast paths (**2, **0.5, **-1, **1) ARE reproduced exactly."""
    import torch
    import finite_difference._native as nat
    rng = np.random.default_rng(77)
    n = 1 << 18
    a = np.abs(rng.standard_normal(n)) * 10.0 ** rng.integers(-3, 4, n) + 1e-6
    b = rng.uniform(-3.0, 3.0, n)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.empty_like(da)
    worst = 0.0
    for s in (1.5, 3.0, -2.5, 0.3333333333333333, 7.0):
        nat.check(nat.lib.fd_ewise_f64(nat.OP_POW_FS, out.data_ptr(), da.data_ptr(), 0, s, n, 0))
        got, want = out.cpu().numpy(), a ** s
        rel = np.max(np.abs(got - want) / np.abs(want))
        worst = max(worst, rel)
        assert rel <= 1e-12, (s, rel)
    nat.check(nat.lib.fd_ewise_f64(nat.OP_POW_FF, out.data_ptr(), da.data_ptr(), db.data_ptr(), 0.0, n, 0))
    got, want = out.cpu().numpy(), a ** b
    rel = np.max(np.abs(got - want) / np.abs(want))
    assert rel <= 1e-12, rel
    assert max(worst, rel) <= 8 * np.finfo(np.float64).eps          # what it really is: a few ulp
    for op, ref in ((nat.OP_SQUARE, a * a), (nat.OP_SQRT, np.sqrt(a)), (nat.OP_RECIP, 1.0 / a), (nat.OP_COPY, a)):
        nat.check(nat.lib.fd_ewise_f64(op, out.data_ptr(), da.data_ptr(), 0, 0.0, n, 0))
        assert np.array_equal(out.cpu().numpy(), ref)


def test_model_run_replay_on_the_device(fd):
    """Model.run(body): the first two steps are recorded, the rest replayed; the
    oracle's bits after 40 steps (lean kernel) and for the staggered pair."""
    shape, n_t = (40, 48, 64), 41
    names = ["x", "y", "z"]
    ic = np.random.default_rng(51).random(shape)
    axes = {"t": np.arange(n_t) * 1e-3}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * 1.0
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=tuple(names), history=20)
        T = fd.Field(m, "T", n_time_ders=1)
        d2 = fd.Stencil([-1, 0, 1], 2)
        T.set_IC_values(ic)
    calls = []

    def body():
        calls.append(1)
        T.dot.assign_update(0.1 * (d2.der(T.prev, "x") + d2.der(T.prev, "y") + d2.der(T.prev, "z")))
        T.time_integrate_update()
    m.run(body)
    want, _ = O.run_heat(shape, n_t, True, 0.1, 1e-3, ic, keep_dot=False)
    assert len(calls) == 2 and m.finished
    for t in (20, 40):
        assert np.array_equal(T.data_level(t), want[t])
    assert m._engine.stats["fused"] == 40 and m._engine.stats["unfused_statements"] == 0
