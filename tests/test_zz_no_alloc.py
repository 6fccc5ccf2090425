"""fd_fused_nested_static_f64: the entry point for hosts that may not allocate
device memory while they step a model (an application that gave the whole GPU
to its own memory pool at start-up, or a time loop that must stay free of
cudaMalloc because that call synchronises the device and can fail late).  Its
scratch level is a fixed array inside the library instead of a caller buffer.

Called here through the C ABI with plain pointers, next to
fd_fused_nested_f64 (caller-provided scratch) on the same inputs: the two must
give the same bits, on two calls in a row -- the array keeps the first call's
contents, and the second call, on a smaller level, must not see them.  Collected
last (the file name sorts after the other tests) and self-contained: it passes
alone in a fresh process and after the rest of the suite.
"""

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _burgers_descriptors(nat, n, u, u_new, flags, scratch_in_descriptor=None):
    """u_new = u + dt * (-0.5 * d/dx(u*u)) on a periodic 1-D grid of n points:
    inner = u*u (expression program), outer = one centred der term of the scratch
    level times -0.5, then the Euler step (FD.py:990-1122, 794-869, 515-516)."""
    inner = nat.Expr()
    ib = inner.base
    ib.n[0], ib.n[1], ib.n[2] = 1, 1, n
    ib.nterms, ib.npost, ib.depth = 0, 0, 1
    ib.dt = 0.0
    ib.flags = flags
    for a in range(3):
        ib.rim_lo[a], ib.rim_hi[a] = 0, max(ib.n[a] - 1, 0)
    inner.nfields, inner.ncode = 1, 2
    inner.field[0] = u
    inner.op[0], inner.arg[0] = nat.X_PUSH_FIELD, 0
    inner.op[1] = nat.X_SQUARE
    # out[0] stays NULL: "the scratch level"

    outer = nat.Fused()
    outer.n[0], outer.n[1], outer.n[2] = 1, 1, n
    outer.nterms, outer.npost, outer.depth = 1, 1, 1
    tm = outer.term[0]
    tm.src = scratch_in_descriptor          # NULL: "the scratch level"
    tm.axis, tm.periodic, tm.lo, tm.hi, tm.n_src = 2, 1, 0, n, n
    tm.st = nat.make_stencil([-1, 0, 1], [-0.5, 0.0, 0.5], 0.5)
    outer.post[0].op, outer.post[0].s = nat.POST_MUL, -0.5
    outer.dt = 1e-2
    outer.prev[0], outer.out[0] = u, u_new
    outer.flags = flags
    for a in range(3):
        outer.rim_lo[a], outer.rim_hi[a] = 0, max(outer.n[a] - 1, 0)
    return inner, outer


def test_static_scratch_entry_point_matches_caller_scratch_twice_in_a_row():
    import torch
    import finite_difference._native as nat
    lib = nat.lib
    flags = torch.zeros(nat.FD_NUM_FLAGS, dtype=torch.int64, device="cuda")
    rng = np.random.default_rng(123)
    for n in (4096, 1000):                   # second call: smaller level, the array still holds the first one's values
        u = torch.from_numpy(1.0 + rng.random(n)).cuda()
        want, got = torch.empty_like(u), torch.empty_like(u)
        scratch = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
        inner, outer = _burgers_descriptors(nat, n, u.data_ptr(), want.data_ptr(), flags.data_ptr())
        nat.check(lib.fd_fused_nested_f64(1, C.addressof(inner), 0, C.addressof(outer), scratch.data_ptr(), n, 0))
        inner, outer = _burgers_descriptors(nat, n, u.data_ptr(), got.data_ptr(), flags.data_ptr())
        nat.check(lib.fd_fused_nested_static_f64(1, C.addressof(inner), 0, C.addressof(outer), 0))
        torch.cuda.synchronize()
        assert torch.equal(got, want)
        # and both are the reference's arithmetic: numpy, operation for operation
        h = u.cpu().numpy()
        sq = h * h
        acc = 0.0 + (-0.5) * np.roll(sq, 1)
        acc = acc + 0.0 * sq
        acc = acc + 0.5 * np.roll(sq, -1)
        ref = h + ((acc / 0.5) * -0.5) * 1e-2
        assert np.array_equal(got.cpu().numpy(), ref)
    assert int(flags[nat.FD_FLAG_INF]) == 0


def test_static_scratch_entry_point_refuses_what_does_not_fit():
    import torch
    import finite_difference._native as nat
    n = nat.FD_STATIC_SCRATCH_BYTES // 8 + 2
    u = torch.zeros(4, dtype=torch.float64, device="cuda")          # never touched: the size test comes first
    inner, outer = _burgers_descriptors(nat, n, u.data_ptr(), u.data_ptr(), 0)
    rc = nat.lib.fd_fused_nested_static_f64(1, C.addressof(inner), 0, C.addressof(outer), 0)
    assert rc != 0
    message = nat.lib.fd_last_error().decode()
    assert "built-in scratch level holds" in message and str(n) in message
    torch.cuda.synchronize()
