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


def test_halo_exchange_one_launch_on_one_gpu():
    """fd_halo_exchange: copy, publication and wait of two fields in ONE launch,
    with this GPU as its own neighbour (periodic ring of one): planes land in the
    ghost planes, the flag words carry the level number, the ticket counter is
    back at zero, and a neighbour that never delivers ends in the time-out flag."""
    import torch
    from finite_difference import _native as nat
    lib = nat.lib
    plane = 64 * 66
    nb = plane * 8
    stream = torch.cuda.current_stream().cuda_stream
    levels = [torch.arange(6 * plane, dtype=torch.float64, device="cuda") + 1000.0 * j for j in range(2)]
    want = [lv.clone() for lv in levels]
    flags = torch.zeros(4, dtype=torch.int64, device="cuda")
    err = torch.zeros(nat.FD_NUM_FLAGS, dtype=torch.int64, device="cuda")
    ticket = torch.zeros(1, dtype=torch.int64, device="cuda")
    links = (nat.HaloLink * 2)()
    for j, lv in enumerate(levels):          # layout [ghost | 4 owned planes | ghost]
        b, f = lv.data_ptr(), flags.data_ptr() + 16 * j
        links[j].src_lo, links[j].peer_dst_lo = b + nb, b + 5 * nb            # lowest owned plane -> high ghost
        links[j].src_hi, links[j].peer_dst_hi = b + 4 * nb, b                 # highest owned plane -> low ghost
        links[j].nbytes = nb
        links[j].peer_flag_lo, links[j].peer_flag_hi = f + 8, f               # "the neighbour's" words are my own
        links[j].my_flag_lo, links[j].my_flag_hi = f, f + 8
        links[j].value = 5
        want[j][5 * plane:] = want[j][plane:2 * plane]
        want[j][:plane] = want[j][4 * plane:5 * plane]
    nat.check(lib.fd_halo_exchange(links, 2, ticket.data_ptr(), 2_000_000_000, err.data_ptr(), stream))
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(levels, want))
    assert flags.tolist() == [5, 5, 5, 5] and int(ticket[0]) == 0 and err.tolist() == [0] * nat.FD_NUM_FLAGS
    # the words this rank waits on are never written: gives up instead of hanging
    lonely = torch.zeros(2, dtype=torch.int64, device="cuda")
    links[0].peer_flag_lo, links[0].peer_flag_hi = lonely.data_ptr(), lonely.data_ptr() + 8
    links[0].value = 6
    nat.check(lib.fd_halo_exchange(links, 1, ticket.data_ptr(), 20_000_000, err.data_ptr(), stream))
    torch.cuda.synchronize()
    assert err.tolist()[nat.FD_FLAG_HALO_TIMEOUT] != 0 and int(ticket[0]) == 0 and lonely.tolist() == [6, 6]
    assert lib.fd_halo_exchange(links, 5, ticket.data_ptr(), 1, err.data_ptr(), stream) != 0


def test_native_slab_step_on_one_gpu_matches_the_oracle():
    """fd_slab_step through the C ABI: a periodic 3-D heat level laid out as ONE
    slab with a ghost plane on either side and this GPU as both neighbours.  The
    interior, the two boundary planes and the exchange of the new level are one
    native call per step; three steps give the single-domain oracle's bits and
    ghost planes that wrap around."""
    import ctypes as C
    import torch
    from finite_difference import _native as nat
    lib = nat.lib
    nx, ny, nz, steps, k, dt = 12, 20, 64, 3, 0.1, 1e-3
    plane = ny * nz
    ic = np.random.default_rng(5).random((nx, ny, nz))
    ring = [torch.full(((nx + 2) * plane,), float("nan"), dtype=torch.float64, device="cuda") for _ in range(2)]
    ring[0][plane:(nx + 1) * plane] = torch.from_numpy(ic.reshape(-1)).cuda()
    ring[0][:plane] = ring[0][nx * plane:(nx + 1) * plane]              # ghosts of the initial level
    ring[0][(nx + 1) * plane:] = ring[0][plane:2 * plane]
    flags = torch.zeros(nat.FD_NUM_FLAGS, dtype=torch.int64, device="cuda")
    words = torch.zeros(4, dtype=torch.int64, device="cuda")             # [slot][low, high]
    counts = torch.zeros(steps + 1, dtype=torch.int64, device="cuda")
    compute = torch.cuda.current_stream()
    comm = torch.cuda.Stream(priority=-1)
    seq = C.c_void_p()
    nat.check(lib.fd_seq_create(C.byref(seq)))
    st = nat.make_stencil([-1, 0, 1], [1.0, -2.0, 1.0], 1.0)
    for t in range(steps):
        src, dst = ring[t % 2], ring[(t + 1) % 2]
        d = nat.Fused()
        d.n[0], d.n[1], d.n[2] = nx + 2, ny, nz
        d.nterms, d.npost, d.depth = 3, 1, 1
        for a in range(3):
            tm = d.term[a]
            tm.src, tm.axis, tm.st = src.data_ptr(), a, st
            tm.periodic, tm.lo, tm.hi, tm.n_src = (0, 1, nx + 1, nx + 2) if a == 0 else (1, 0, d.n[a], d.n[a])
            d.box_lo[a], d.box_hi[a] = 0, d.n[a]
            d.rim_lo[a], d.rim_hi[a] = 0, d.n[a] - 1
        d.rim_lo[0], d.rim_hi[0] = 1, nx
        d.post[0].op, d.post[0].s = nat.POST_MUL, k
        d.dt = dt
        d.prev[0], d.out[0] = src.data_ptr(), dst.data_ptr()
        d.nan_count[0] = counts.data_ptr() + 8 * (t + 1)
        d.flags = flags.data_ptr()
        step = nat.SlabStep()
        step.kind, step.dim, step.lo, step.hi, step.depth, step.nlinks = 0, 0, 1, nx + 1, 1, 1
        step.desc = C.addressof(d)
        b, f = dst.data_ptr(), words.data_ptr() + 16 * ((t + 1) % 2)
        ln = step.link[0]
        ln.src_lo, ln.peer_dst_lo = b + plane * 8, b + (nx + 1) * plane * 8
        ln.src_hi, ln.peer_dst_hi = b + nx * plane * 8, b
        ln.nbytes = plane * 8
        ln.peer_flag_lo, ln.peer_flag_hi, ln.my_flag_lo, ln.my_flag_hi = f + 8, f, f, f + 8
        ln.value = t + 2
        step.timeout_ns, step.flags = 2_000_000_000, flags.data_ptr()
        step.compute_stream, step.comm_stream = compute.cuda_stream, comm.cuda_stream
        nat.check(lib.fd_slab_step_f64(seq, C.byref(step)))
        nat.check(lib.fd_seq_wait_halo(seq, compute.cuda_stream))       # the next step reads the ghosts
    torch.cuda.synchronize()
    nat.check(lib.fd_seq_destroy(seq))
    want, _ = O.run_heat((nx, ny, nz), steps + 1, True, k, dt, ic, keep_dot=False)
    got = ring[steps % 2].cpu().numpy().reshape(nx + 2, ny, nz)
    assert np.array_equal(got[1:nx + 1], want[steps])
    assert np.array_equal(got[0], want[steps][-1]) and np.array_equal(got[nx + 1], want[steps][0])
    assert flags.tolist() == [0] * nat.FD_NUM_FLAGS and counts.tolist() == [0] * (steps + 1)
    assert words.tolist() == [steps, steps, steps + 1, steps + 1] or words.tolist() == [steps + 1, steps + 1, steps, steps]
