"""Host-side logic of the package on the CPU-only container: the same
scenarios as the GPU parity tests, with libfd_b200.so and the device replaced
by the numpy stand-ins of tests/fake_native.py (test infrastructure)."""

import pytest

import finite_difference as fd
from oracle import scenarios
from tests import fake_native
from tests.scenario_check import run_and_compare


@pytest.mark.parametrize("name", sorted(scenarios.ALL))
def test_scenario_fake_native(name):
    with fake_native.installed() as fake:
        run_and_compare(name, fd)
        assert fake.launches > 0


def test_staggered_pair_is_one_launch_per_step():
    import finite_difference.model as M
    engines = []
    orig = M.Model._get_engine

    def spy(self):
        e = orig(self)
        if e not in engines:
            engines.append(e)
        return e

    M.Model._get_engine = spy
    try:
        with fake_native.installed():
            run_and_compare("acoustic3d_wide", fd)
            assert engines[-1].stats.get("fused_pairs") == 4 and engines[-1].stats["fused"] == 0
            run_and_compare("acoustic2d", fd)
            assert engines[-1].stats.get("fused_pairs") == 4
    finally:
        M.Model._get_engine = orig


@pytest.mark.parametrize("name", ["heat2d_fixed", "wave2d", "acoustic3d_fixed", "flux_1d"])
def test_scenario_unfused_only(name, monkeypatch):
    monkeypatch.setenv("FD_B200_NO_FUSE", "1")
    with fake_native.installed():
        run_and_compare(name, fd)


@pytest.mark.parametrize("shape", [(10, 400), (16, 30, 768)])
def test_interior_rim_split_fake_native(shape):
    """Fixed boundaries on a grid wide enough for the interior/rim split: the
    statement is issued as one predicate-free interior box plus rim slabs, and
    the result is still the oracle's, bit for bit."""
    import contextlib
    import io
    import numpy as np
    from oracle import fd_numpy as O
    names = ["x", "y", "z"][:len(shape)]
    ic = np.random.default_rng(21).random(shape)
    with fake_native.installed() as fake:
        axes = {"t": np.arange(4) * 1e-3}
        for n, a in zip(shape, names):
            axes[a] = np.arange(n) * 1.0
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model(axes, "t")
            T = fd.Field(m, "T", n_time_ders=1)
            d2 = fd.Stencil([-1, 0, 1], 2)
            T.set_IC_values(ic)
            for a in names:
                T.set_BC("0", a, "start")
                T.set_BC("1", a, "end")
            while not m.finished:
                acc = d2.der(T.prev, names[0])
                for a in names[1:]:
                    acc = acc + d2.der(T.prev, a)
                T.dot.assign_update(0.1 * acc)
                T.time_integrate_update()
                m.increment_time()
        got, got_dot = T.data, T.dot.data
        # interior box + the rim slabs of the non-contiguous axes per step (3-point
        # stencils: the contiguous axis' rim is handled inside the kernels)
        assert fake.launches >= 3 * (1 + 2 * (len(shape) - 1))
    # oracle with the same boundary values
    g = O.Grid(axes, "t")
    S, Sd = O.Series(g), O.Series(g)
    S.data[0] = ic
    for a in names:
        S.set_face(a, "start", 0.0)
        S.set_face(a, "end", 1.0)
    st = O.StencilSpec([-1, 0, 1], 2)
    while g.step + 1 < g.n_t:
        acc = st.apply(g, S.data[g.step], names[0])
        for a in names[1:]:
            acc = acc + st.apply(g, S.data[g.step], a)
        Sd.store(g.step + 1, 0.1 * acc)
        O.integrate(S, Sd, g)
        g.step += 1
    assert np.array_equal(got, S.data, equal_nan=True)
    assert np.array_equal(got_dot, Sd.data, equal_nan=True)


def test_time_varying_coefficient_does_not_reuse_a_stale_plan():
    """The launch descriptor is cached per statement; a scalar that changes from
    step to step is part of the statement's key."""
    import contextlib
    import io
    import numpy as np
    from oracle import fd_numpy as O
    n, n_t = 16, 6
    ic = np.random.default_rng(5).random(n)
    with fake_native.installed():
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model({"t": np.arange(n_t) * 0.01, "x": np.arange(n) * 1.0}, "t", periodic=("x",))
            T = fd.Field(m, "T", n_time_ders=1)
            d2 = fd.Stencil([-1, 0, 1], 2)
            T.set_IC_values(ic)
            while not m.finished:
                T.dot.assign_update((0.1 + 0.05 * m.time_step) * d2.der(T.prev, "x"))
                T.time_integrate_update()
                m.increment_time()
        got = T.data
    want = np.full((n_t, n), np.nan)
    want[0] = ic
    w = O.stencil_weights([-1, 0, 1], 2)
    for t in range(n_t - 1):
        lap = O.der(want[t], [-1, 0, 1], w, 2, 0, True, False, False, 1.0)
        want[t + 1] = want[t] + ((0.1 + 0.05 * t) * lap) * 0.01
    assert np.array_equal(got, want)


def test_peer_links_address_arithmetic():
    """PeerLinks (peer-memory halo exchange) with stand-ins for the C ABI and the
    process group: slabs of unequal thickness, fixed and periodic slab axes, and
    world 2 where both neighbours are the same rank.  Checked: which planes are
    sent where, which flag words are written and awaited."""
    import ctypes as C
    from finite_difference import _dist

    class T:                                   # a 'device tensor'
        def __init__(self, addr):
            self.addr = addr

        def data_ptr(self):
            return self.addr

    class Lib:                                 # allocation r*2^32 holds everything rank r exports
        def __init__(self, rank, calls):
            self.rank, self.calls = rank, calls

        def fd_ipc_export(self, ptr, handle, off):
            handle.raw = bytes([ptr >> 32]) * 64
            off._obj.value = ptr & 0xffffffff
            return 0

        def fd_ipc_open(self, handle, base):
            owner = handle[0]
            assert owner != self.rank
            base._obj.value = (owner << 32) + 0x7000_0000_0000       # mapped somewhere else in this process
            return 0

        def fd_halo_push(self, *a):
            self.calls.append(("push",) + a)
            return 0

        def fd_halo_wait(self, *a):
            self.calls.append(("wait",) + a)
            return 0

    class Nat:
        FD_IPC_HANDLE_BYTES = 64
        NativeError = RuntimeError

        def __init__(self, rank, calls):
            self.lib = Lib(rank, calls)

        @staticmethod
        def check(rc):
            assert rc == 0

    class Dist:
        def __init__(self, world, table):
            self.world, self.table = world, table

        def all_gather_object(self, out, mine):
            out[:] = self.table.pop(0)(mine)

    plane_bytes, H = 4096, 2
    for world, periodic, counts in [(3, False, [5, 4, 4]), (3, True, [5, 4, 4]), (2, True, [6, 6])]:
        ring_addr = lambda r, k: (r << 32) + 0x10000 * (k + 1)
        flag_addr = lambda r: (r << 32) + 0x80000
        exports = []
        for r in range(world):                                     # what every rank would publish
            exports.append({"ring": [(bytes([r]) * 64, ring_addr(r, k) & 0xffffffff) for k in range(2)],
                            "flags": (bytes([r]) * 64, flag_addr(r) & 0xffffffff), "owned0": counts[r]})
        for r in range(world):
            _dist.PeerLinks._opened = {}
            calls = []

            class Slab:
                pass
            slab = Slab()
            slab.world, slab.rank = world, r
            slab.prev_rank = (r - 1) % world if (periodic or r > 0) else None
            slab.next_rank = (r + 1) % world if (periodic or r < world - 1) else None

            def gather_entries(mine, r=r):
                assert mine == exports[r]
                return list(exports)
            slab.dist = Dist(world, [gather_entries, lambda ok: [ok] * world])
            links = _dist.PeerLinks(slab, Nat(r, calls), [T(ring_addr(r, 0)), T(ring_addr(r, 1))], T(flag_addr(r)),
                                    plane_bytes, counts[r], H)
            assert links.ok and links.nbytes == H * plane_bytes
            mapped = lambda q, addr: addr + 0x7000_0000_0000
            for k in range(2):
                level = 11 + links.base             # (the second round comes after a retired model)
                links.push(k, 11, 0)
                links.wait(k, 11, 123, 0)
                push, wait = calls[-2], calls[-1]
                prev, nxt = slab.prev_rank, slab.next_rank
                base = ring_addr(r, k)
                want_push = (
                    base + H * plane_bytes if prev is not None else None,                                   # my low owned planes
                    mapped(prev, ring_addr(prev, k)) + (H + counts[prev]) * plane_bytes if prev is not None else None,  # prev's high ghosts
                    base + counts[r] * plane_bytes if nxt is not None else None,                          # my high owned planes
                    mapped(nxt, ring_addr(nxt, k)) if nxt is not None else None,                          # next's low ghosts
                    H * plane_bytes,
                    mapped(prev, flag_addr(prev)) + (2 * k + 1) * 8 if prev is not None else None,        # "your high ghost is fresh"
                    mapped(nxt, flag_addr(nxt)) + (2 * k) * 8 if nxt is not None else None,               # "your low ghost is fresh"
                    level, 0)
                assert push[1:] == want_push
                assert wait[1:] == (flag_addr(r) + 2 * k * 8 if prev is not None else None,
                                    flag_addr(r) + (2 * k + 1) * 8 if nxt is not None else None, level, links.TIMEOUT_NS, 123, 0)
                # the same exchange as one fd_halo_link_t (fd_halo_exchange / fd_slab_step), after a
                # retired model: level numbers continue above everything the old model published
                import finite_difference._native as real_nat
                link = real_nat.HaloLink()
                links.retire(40)
                links.fill(link, k, 11)
                assert (link.src_lo, link.peer_dst_lo, link.src_hi, link.peer_dst_hi, link.nbytes, link.peer_flag_lo,
                        link.peer_flag_hi) == want_push[:7]
                assert (link.my_flag_lo, link.my_flag_hi) == wait[1:3] and link.value == 11 + links.base
                assert links.base == 42 * (k + 1)
    # planes that are not multiples of 16 bytes: the field keeps isend/irecv, on every rank
    slab = Slab()
    slab.world, slab.rank, slab.prev_rank, slab.next_rank = 2, 0, 1, 1
    slab.dist = Dist(2, [lambda mine: [mine, {"ring": [], "flags": None, "owned0": 4}], lambda ok: [ok, True]])
    assert not _dist.PeerLinks(slab, Nat(0, []), [T(0x10000)], T(0x80000), 360, 4, 1).ok


def test_auxiliary_assign_is_fused():
    """An assign_update that no time_integrate_update follows (auxiliary flux
    field with boundary faces) is one fused launch: no statement
    of the flux scenario takes the operator-by-operator path."""
    import finite_difference.model as M
    engines = []
    orig = M.Model._get_engine

    def spy(self):
        e = orig(self)
        if e not in engines:
            engines.append(e)
        return e

    M.Model._get_engine = spy
    try:
        with fake_native.installed():
            run_and_compare("flux_1d", fd)
    finally:
        M.Model._get_engine = orig
    assert engines and all(e.stats["unfused_statements"] == 0 and e.stats["fused"] == 10 for e in engines)


def test_more_than_three_spatial_axes_host_planner():
    """CPU twin of the GPU test of the same name: the planner folds the leading
    axes and hands the stand-in the same descriptors the kernels get."""
    import numpy as np
    from tests.nd_cases import CASES, run_case
    with fake_native.installed():
        for shape, periodic, n_ders in CASES:
            got, want, stats = run_case(fd, shape, periodic, n_ders)
            assert np.array_equal(got, want, equal_nan=True), (shape, periodic)
            if periodic:
                assert stats["fused"] == 3 and stats["unfused_statements"] == 0, (shape, stats)


def test_constant_field_leaf_keeps_its_plan():
    """A coefficient field (ConstantField.always) is one fixed level: the plan of
    a statement that uses it is found again every step instead of being rebuilt."""
    import contextlib
    import io
    import numpy as np
    with fake_native.installed():
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model({"t": np.arange(9) * 1e-2, "x": np.arange(10) * 1.0, "y": np.arange(8) * 1.0}, "t",
                         periodic=("x", "y"))
            T = fd.Field(m, "T", n_time_ders=1)
            kappa = fd.ConstantField(m, "kappa")
            kappa.set("1+x*y/50")
            d2 = fd.Stencil([-1, 0, 1], 2)
            T.set_IC("x+y")
            while not m.finished:
                T.dot.assign_update(kappa.always * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
                T.time_integrate_update()
                m.increment_time()
        assert m._engine.stats["fused"] == 8
        assert len(m._engine._plans) == 2          # one per ring parity


@pytest.mark.parametrize("name,launches", [("nested_burgers", 2), ("nested_staggered_flux", 2), ("nested_divgrad", 3)])
def test_der_of_a_computed_expression_is_fused(name, launches):
    """One scratch level per inner expression plus the statement itself: 2 (or
    3) launches per step instead of one per reference array operation."""
    made = []
    real = fd.Model

    def spy(*a, **k):
        made.append(real(*a, **k))
        return made[-1]
    with fake_native.installed() as lib:
        fd.Model = spy
        try:
            run_and_compare(name, fd, check_prints=False)
        finally:
            fd.Model = real
        stats = made[-1]._engine.stats
        assert stats["unfused_statements"] == 0 and stats["fused_nested"] == 4, stats
        assert lib.launches - 4 * launches in (1, 2), lib.launches   # + NaN count (and device fill) of the initial level


def test_model_close_keeps_results_and_refuses_further_steps():
    """Model.close() (additive): the history is brought up to date first, the
    device state goes, and a closed model cannot be advanced."""
    import contextlib
    import io
    import numpy as np
    from oracle import fd_numpy as O
    shape, n_t = (6, 8), 5
    ic = np.random.default_rng(9).random(shape)
    with fake_native.installed():
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model({"t": np.arange(n_t) * 1e-3, "x": np.arange(shape[0]) * 1.0, "y": np.arange(shape[1]) * 1.0}, "t",
                         periodic=("x", "y"))
            T = fd.Field(m, "T", n_time_ders=1)
            d2 = fd.Stencil([-1, 0, 1], 2)
            T.set_IC_values(ic)
            for _ in range(3):
                T.dot.assign_update(0.1 * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
                T.time_integrate_update()
                m.increment_time()
            m.close()
            m.close()                                    # idempotent
            want, _ = O.run_heat(shape, n_t, True, 0.1, 1e-3, ic)
            assert np.array_equal(T.data[:4], want[:4], equal_nan=True)
            assert m._engine.states == {}
            T.dot.assign_update(0.1 * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
            T.time_integrate_update()
            with pytest.raises(Exception, match="closed"):
                m.increment_time()


def _models_for_run():
    """(name, builder) pairs: builder() -> (model, body, fields to compare)."""
    import contextlib
    import io
    import numpy as np

    def heat_fixed_time_bc():
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model({"t": np.arange(9) * 1e-2, "x": np.arange(14) * 1.0, "y": np.arange(16) * 1.0}, "t")
            T = fd.Field(m, "T", n_time_ders=1)
            d2 = fd.Stencil([-1, 0, 1], 2)
            T.set_IC("x+y")
            T.set_BC("0", "x", "start")
            T.set_BC("t*y", "x", "end")          # boundary values that change with time
            T.set_BC("1", "y", "start")
            T.set_BC("2", "y", "end")

        def body():
            T.dot.assign_update(0.1 * (d2.der(T.prev, "x") + d2.der(T.prev, "y")))
            T.time_integrate_update()
        return m, body, [T, T.dot]

    def wave():
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model({"t": np.arange(9) * 1e-2, "x": np.arange(16) * 1.0, "y": np.arange(16) * 1.0}, "t",
                         periodic=("x", "y"))
            u = fd.Field(m, "u", n_time_ders=2)
            st = fd.Stencil(list(range(-2, 3)), 2)
            u.set_IC("x*y")
            u.dot.set_IC("0")

        def body():
            u.dot.dot.assign_update(0.01 * (st.der(u.prev, "x") + st.der(u.prev, "y")))
            u.dot.time_integrate_update()
            u.time_integrate_update()
        return m, body, [u, u.dot]

    def acoustic():
        with contextlib.redirect_stdout(io.StringIO()):
            names = ("x", "y", "z")
            m = fd.Model({"t": np.arange(9) * 0.1, "x": np.arange(6) * 1.0, "y": np.arange(6) * 1.0, "z": np.arange(8) * 1.0},
                         "t", periodic=names)
            p = fd.Field(m, "p", n_time_ders=1)
            v = {a: fd.Field(m, "v" + a, edge_axes=(a,), n_time_ders=1) for a in names}
            ce, ec = fd.Stencil([-.5, .5], 1, "cell", "edge"), fd.Stencil([-.5, .5], 1, "edge", "cell")
            p.set_IC("x*y+z")
            for a in names:
                v[a].set_IC("0")

        def body():
            for a in names:
                v[a].dot.assign_update(-1.0 * ce.der(p.prev, a))
                v[a].time_integrate_update()
            p.dot.assign_update(-0.25 * (ec.der(v["x"].new, "x") + ec.der(v["y"].new, "y") + ec.der(v["z"].new, "z")))
            p.time_integrate_update()
        return m, body, [p, v["x"], v["z"]]

    def nested():
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model({"t": np.arange(9) * 1e-2, "x": np.arange(40) * 0.5}, "t")
            u = fd.Field(m, "u", n_time_ders=1)
            d1, d2 = fd.Stencil([-1, 0, 1], 1), fd.Stencil([-1, 0, 1], 2)
            u.set_IC("1+x/40")
            u.set_BC("1", "x", "start")
            u.set_BC("1.5", "x", "end")

        def body():
            u.dot.assign_update(-0.5 * d1.der(u.prev * u.prev, "x") + 0.05 * d2.der(u.prev, "x"))
            u.time_integrate_update()
        return m, body, [u, u.dot]

    return [("heat_fixed_time_bc", heat_fixed_time_bc), ("wave", wave), ("acoustic", acoustic), ("nested", nested)]


@pytest.mark.parametrize("unfused", [False, True])
def test_model_run_replays_the_recorded_step(unfused, monkeypatch):
    """Model.run(body) (additive): body is called for the first two steps only,
    afterwards the recorded statements are replayed with their time levels moved
    on -- same arrays as the written-out loop, with the history policy, boundary
    values that depend on time, the staggered pair, a nested der, and the unfused
    operators; also run(body, k) in pieces mixed with the written-out loop."""
    import numpy as np
    if unfused:
        monkeypatch.setenv("FD_B200_NO_FUSE", "1")
    with fake_native.installed():
        for name, build in _models_for_run():
            m, body, fields = build()
            while not m.finished:
                body()
                m.increment_time()
            want = [np.array(f.data) for f in fields]
            m, body, fields = build()
            calls = []
            m.run(lambda: (calls.append(1), body()))
            assert m.finished and len(calls) == 2, (name, len(calls))
            for f, w in zip(fields, want):
                assert np.array_equal(f.data, w, equal_nan=True), (name, f.name)
            m, body, fields = build()
            m.run(body, 3)
            body()
            m.increment_time()
            m.run(body, 2)
            m.run(body)
            assert m.finished
            for f, w in zip(fields, want):
                assert np.array_equal(f.data, w, equal_nan=True), (name, f.name, "pieces")
