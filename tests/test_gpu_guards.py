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


def test_halo_push_and_wait_on_one_gpu():
    """fd_halo_push / fd_halo_wait through the C ABI with both 'peers' on this
    GPU: planes copied, level number published after the copy, the wait returns
    once both flag words have reached it, and gives up with
    FD_FLAG_HALO_TIMEOUT (instead of hanging) when a neighbour never delivers."""
    import ctypes as C
    import torch
    from finite_difference import _native as nat
    lib = nat.lib
    plane = 96 * 130
    src = torch.arange(4 * plane, dtype=torch.float64, device="cuda")
    dst = torch.full((4 * plane,), -1.0, dtype=torch.float64, device="cuda")
    flags = torch.zeros(4, dtype=torch.int64, device="cuda")
    err = torch.zeros(nat.FD_NUM_FLAGS, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    base_s, base_d, f = src.data_ptr(), dst.data_ptr(), flags.data_ptr()
    nbytes = plane * 8
    # low side: plane 1 -> plane 3 of dst; high side: plane 2 -> plane 0 of dst
    nat.check(lib.fd_halo_push(base_s + nbytes, base_d + 3 * nbytes, base_s + 2 * nbytes, base_d, nbytes, f, f + 8, 7, stream))
    nat.check(lib.fd_halo_wait(f, f + 8, 7, 2_000_000_000, err.data_ptr(), stream))
    torch.cuda.synchronize()
    assert flags.tolist() == [7, 7, 0, 0] and err.tolist() == [0] * nat.FD_NUM_FLAGS
    assert torch.equal(dst[3 * plane:], src[plane:2 * plane]) and torch.equal(dst[:plane], src[2 * plane:3 * plane])
    assert bool((dst[plane:3 * plane] == -1.0).all())
    # one-sided (end rank of a fixed slab axis): only the high side exists
    nat.check(lib.fd_halo_push(None, None, base_s, base_d + nbytes, nbytes, None, f + 16, 9, stream))
    nat.check(lib.fd_halo_wait(None, f + 16, 9, 2_000_000_000, err.data_ptr(), stream))
    torch.cuda.synchronize()
    assert flags.tolist() == [7, 7, 9, 0] and torch.equal(dst[plane:2 * plane], src[:plane])
    # an older level number is already satisfied; a newer one that never comes times out
    nat.check(lib.fd_halo_wait(f, f + 8, 5, 2_000_000_000, err.data_ptr(), stream))
    nat.check(lib.fd_halo_wait(f, f + 24, 8, 20_000_000, err.data_ptr(), stream))
    torch.cuda.synchronize()
    assert err.tolist()[nat.FD_FLAG_HALO_TIMEOUT] != 0 and err.tolist()[nat.FD_FLAG_INF] == 0
    # argument checks
    assert lib.fd_halo_push(base_s, base_d, None, None, 24, None, None, 1, stream) != 0
    assert b"multiple of 16" in lib.fd_last_error()
    assert lib.fd_halo_push(base_s, None, None, None, 32, None, None, 1, stream) != 0
