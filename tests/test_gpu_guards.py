"""Out-of-bounds check without a sanitizer: every device buffer the engine
allocates is embedded between two guard zones filled with a huge finite
sentinel.  A kernel that reads outside its buffer pulls the sentinel into the
result (mismatch with the oracle); one that writes outside changes a guard."""

import contextlib
import io

import numpy as np
import pytest

from oracle import fd_numpy as O

pytestmark = pytest.mark.gpu

PAD = 2048
SENTINEL = 7.25e77


@contextlib.contextmanager
def guarded_allocations():
    import torch
    from finite_difference._device import Device, _torch_dtype
    bases = []
    orig = Device.empty

    def empty(self, n, dtype):
        sentinel = SENTINEL if np.dtype(dtype) == np.float64 else 7.25e30
        base = torch.full((int(n) + 2 * PAD,), sentinel, dtype=_torch_dtype(dtype), device=self.dev)
        bases.append((base, int(n), sentinel))
        return base[PAD:PAD + int(n)]

    Device.empty = empty
    try:
        yield bases
    finally:
        Device.empty = orig


def _check_guards(bases):
    import torch
    torch.cuda.synchronize()
    for base, n, sentinel in bases:
        lo, hi = base[:PAD], base[PAD + n:]
        assert bool((lo == sentinel).all()) and bool((hi == sentinel).all()), "a kernel wrote outside its buffer"


SHAPES = [((37,), True), ((64,), False), ((130, 66), True), ((33, 258), False), ((20, 24, 64), True),
          ((21, 23, 130), True), ((19, 22, 66), False), ((12, 31, 250), True), ((9, 10, 512), False)]


@pytest.mark.parametrize("shape,periodic", SHAPES)
@pytest.mark.parametrize("points", [(-1, 0, 1), (-2, -1, 0, 1, 2)])
def test_heat_with_guard_zones(shape, periodic, points):
    import finite_difference as fd
    if min(shape) < len(points) + 2:
        pytest.skip("axis too short for this stencil")
    names = ["x", "y", "z"][:len(shape)]
    ic = np.random.default_rng(61).random(shape)
    with guarded_allocations() as bases, contextlib.redirect_stdout(io.StringIO()):
        axes = {"t": np.arange(4) * 1e-3}
        for n, a in zip(shape, names):
            axes[a] = np.arange(n) * 1.0
        m = fd.Model(axes, "t", periodic=tuple(names) if periodic else (), history="last")
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
            T.dot.assign_update(0.1 * acc)
            T.time_integrate_update()
            m.increment_time()
        got = T.data_level(3)
        _check_guards(bases)
    want, _ = O.run_heat(shape, 4, periodic, 0.1, 1e-3, ic, stencil_points=points)
    assert np.array_equal(got, want[3], equal_nan=True)


@pytest.mark.parametrize("march", [False, True])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_staggered_pair_with_guard_zones(dtype, march, monkeypatch):
    import finite_difference as fd
    from tests.test_gpu_parity import _acoustic
    if march:   # a forced chunk length also forces the marching kernel on this small grid
        monkeypatch.setenv("FD_B200_PAIR_CHUNK", "4")
    shape, n_t = (10, 12, 40), 5
    ic = np.random.default_rng(62).random(shape).astype(dtype)
    with guarded_allocations() as bases:
        m, p, v = _acoustic(fd, shape, n_t, True, ic, dtype)
        got = p.data
        _check_guards(bases)
    want, _ = O.run_acoustic(shape, n_t, 1.0, 0.25, 0.1, ic, dtype=dtype)
    assert np.array_equal(got, want.astype(np.float64), equal_nan=True)
    assert m._engine.stats.get("fused_pairs") == n_t - 1
