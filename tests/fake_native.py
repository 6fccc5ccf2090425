"""TEST INFRASTRUCTURE ONLY: a numpy stand-in for ``libfd_b200.so`` and for the
CUDA device, so that the *host-side* logic of the package (deferred DAG,
pattern matcher, ring buffer bookkeeping, boundary-face extraction, history)
can be exercised in the CPU-only container.  It is never imported by the
package; the ``-m gpu`` tests run the same scenarios through the real library.

"Device memory" is CPU torch tensors; the entry points read their arguments
exactly as the C ABI declares them (addresses + sizes + descriptor structs)
and compute with the oracle's numpy operations.
"""

from __future__ import annotations

import ctypes as C
import contextlib

import numpy as np
import torch

from oracle import fd_numpy as O


def _arr(address, n, dtype=np.float64):
    if n == 0:
        return np.empty(0, dtype=dtype)
    ctype = {np.dtype(np.float64): C.c_double, np.dtype(np.float32): C.c_float,
             np.dtype(np.int64): C.c_int64}[np.dtype(dtype)]
    return np.ctypeslib.as_array((ctype * int(n)).from_address(int(address)))


def _val(p):
    return p.value if hasattr(p, "value") else p


class FakeLib:
    """Subset of the C ABI in numpy, float64 only."""

    def __init__(self):
        self.launches = 0
        self.error = b""

    def fd_abi_version(self):
        return 6

    def fd_last_kernel(self):
        return b"numpy stand-in"

    def fd_last_error(self):
        return self.error

    def fd_launch_count(self):
        return self.launches

    # -- helpers ------------------------------------------------------------
    @staticmethod
    def _stencil(st):
        n = st.npts
        return list(st.shifts[:n]), list(st.weights[:n]), st.scale

    @staticmethod
    def _der(src, axis, shifts, weights, scale, periodic, lo, hi, n_dst, flags):
        if periodic:
            out = np.zeros(src.shape)
            for w, s in zip(weights, shifts):
                out += w * np.roll(src, -s, axis=axis)
        else:
            shape = list(src.shape)
            shape[axis] = n_dst
            out = np.full(shape, np.nan)
            part_shape = list(shape)
            part_shape[axis] = hi - lo
            part = np.zeros(part_shape)
            sel = [slice(None)] * src.ndim
            for w, s in zip(weights, shifts):
                sel[axis] = slice(lo + s, hi + s)
                part += w * src[tuple(sel)]
            sel[axis] = slice(lo, hi)
            out[tuple(sel)] = part
        out = out / scale
        if flags is not None and np.any(np.isinf(out)):
            flags[0] = 1
        return out

    # -- unfused ------------------------------------------------------------
    def fd_der_f64(self, dst, src, outer, n_src, n_dst, inner, st, periodic, lo, hi, flags, stream):
        self.launches += 1
        st = st._obj if hasattr(st, "_obj") else st.contents
        shifts, weights, scale = self._stencil(st)
        s = _arr(src, outer * n_src * inner).reshape(outer, n_src, inner)
        f = _arr(flags, 4, np.int64) if flags else None
        out = self._der(s, 1, shifts, weights, scale, periodic, lo, hi, n_dst, f)
        _arr(dst, outer * n_dst * inner)[:] = out.reshape(-1)
        return 0

    def fd_ewise_f64(self, op, dst, a, b, s, n, stream):
        self.launches += 1
        x = np.array(_arr(a, n))
        y = np.array(_arr(b, n)) if b else None
        with np.errstate(all="ignore"):
            r = {0: lambda: -x, 1: lambda: x + y, 2: lambda: x - y, 3: lambda: x * y, 4: lambda: x / y,
                 5: lambda: x ** y, 6: lambda: x + s, 7: lambda: x - s, 8: lambda: x * s, 9: lambda: x / s,
                 10: lambda: x ** s, 11: lambda: x * x, 12: lambda: 1 / x, 13: lambda: np.sqrt(x),
                 14: lambda: x, 15: lambda: x + y * s}[op]()
        _arr(dst, n)[:] = r
        return 0

    def fd_merge_store_f64(self, stored, fresh, n, nan_count, flags, stream):
        self.launches += 1
        st, fr = _arr(stored, n), _arr(fresh, n)
        known = ~np.isnan(fr)
        if np.any(~np.isnan(st[known])) and flags:
            _arr(flags, 4, np.int64)[1] = 1
        st[known] = fr[known]
        if nan_count:
            _arr(nan_count, 1, np.int64)[0] += int(np.isnan(st).sum())
        return 0

    def fd_fill_f64(self, dst, value, n, stream):
        self.launches += 1
        _arr(dst, n)[:] = value
        return 0

    def fd_face_store_f64(self, stored, face, outer, n, inner, index, stream):
        self.launches += 1
        st = _arr(stored, outer * n * inner).reshape(outer, n, inner)
        fc = _arr(face, outer * inner).reshape(outer, inner)
        known = ~np.isnan(fc)
        st[:, index, :][known] = fc[known]
        return 0

    def fd_count_nan_f64(self, src, n, nan_count, stream):
        self.launches += 1
        _arr(nan_count, 1, np.int64)[0] += int(np.isnan(_arr(src, n)).sum())
        return 0

    # -- fused --------------------------------------------------------------
    def _term(self, tm, n, flags):
        """One der term of a fused descriptor on the whole (n0, n1, n2) view."""
        shifts, weights, scale = self._stencil(tm.st)
        if tm.fold_mod and not (tm.fold_div == 1 and tm.fold_mod == n[0]):
            # dimension 0 merges several leading axes; differentiate the folded one
            view = (n[0] // (tm.fold_div * tm.fold_mod), tm.fold_mod, tm.fold_div, n[1], n[2])
            src = _arr(tm.src, int(np.prod(n))).reshape(view)
            return self._der(src, 1, shifts, weights, scale, tm.periodic, tm.lo, tm.hi, tm.fold_mod, flags).reshape(n)
        sshape = list(n)
        sshape[tm.axis] = tm.n_src
        src = _arr(tm.src, int(np.prod(sshape))).reshape(sshape)
        return self._der(src, tm.axis, shifts, weights, scale, tm.periodic, tm.lo, tm.hi, n[tm.axis], flags)

    def fd_fused_step_f64(self, desc, stream):
        self.launches += 1
        d = desc._obj if hasattr(desc, "_obj") else desc.contents
        n = tuple(d.n)
        size = int(np.prod(n))
        flags = _arr(d.flags, 4, np.int64) if d.flags else None
        total = None
        for k in range(d.nterms):
            tm = d.term[k]
            v = self._term(tm, n, flags)
            if tm.negate:
                v = -v
            if tm.has_pre:
                v = v * tm.pre
            total = v if k == 0 else (total - v if tm.subtract else total + v)
        for k in range(d.npost):
            po = d.post[k]
            if po.op == 0:
                total = total * po.s
            elif po.op == 1:
                total = total + po.s
            elif po.op == 2:
                total = total - po.s
            elif po.op == 3:
                total = total / po.s
            elif po.op == 4:
                total = -total
            elif po.op == 5:
                total = total * _arr(po.coef, size).reshape(n)
        if any(d.box_hi):
            region = tuple(slice(d.box_lo[a], d.box_hi[a]) for a in range(3))
        else:
            region = (slice(None),) * 3
        if d.dot_out:
            _arr(d.dot_out, size).reshape(n)[region] = total[region]
        rate = total
        for l in range(d.depth):
            new = _arr(d.prev[l], size).reshape(n) + rate * d.dt if d.prev[l] else np.array(rate, dtype=float)
            stored = np.full(n, np.nan)
            for dim in range(3):
                for side in range(2):
                    fp = d.face[l][dim][side]
                    if not fp:
                        continue
                    fshape = [n[a] for a in range(3) if a != dim]
                    face = _arr(fp, int(np.prod(fshape))).reshape(fshape)
                    sel = [slice(None)] * 3
                    sel[dim] = d.rim_lo[dim] if side == 0 else d.rim_hi[dim]
                    tgt = stored[tuple(sel)]
                    fill = np.isnan(tgt)
                    tgt[fill] = face[fill]
            known = ~np.isnan(new)
            if flags is not None and np.any(~np.isnan(stored[region][known[region]])):
                flags[1] = 1
            new[~known] = stored[~known]
            out = _arr(d.out[l], size).reshape(n)
            out[region] = new[region]
            if d.nan_count[l]:
                _arr(d.nan_count[l], 1, np.int64)[0] += int(np.isnan(new[region]).sum())
            rate = new
        return 0


def _pair(self, desc, stream):
    """fd_fused_pair_f64 in numpy (periodic rolls; the box keeps clear of wrapped ghosts)."""
    self.launches += 1
    d = desc._obj if hasattr(desc, "_obj") else desc.contents
    n = tuple(d.n)
    size = int(np.prod(n))
    region = tuple(slice(d.box_lo[a], d.box_hi[a]) for a in range(3)) if any(d.box_hi) else (slice(None),) * 3
    flags = _arr(d.flags, 4, np.int64) if d.flags else None
    p = _arr(d.p_prev, size).reshape(n)
    total = None
    for k in range(d.naxes):
        A = d.ax[k]
        u, w = list(A.up.weights[:2]), list(A.down.weights[:2])
        v = _arr(A.v_prev, size).reshape(n)
        g = ((0.0 + u[0] * np.roll(p, 1, axis=A.axis)) + u[1] * p) / A.up.scale
        rhs = g * A.c_up
        vnew = v + rhs * d.dt
        dd = ((0.0 + w[0] * vnew) + w[1] * np.roll(vnew, -1, axis=A.axis)) / A.down.scale
        if flags is not None and (np.any(np.isinf(g[region])) or np.any(np.isinf(dd[region]))):
            flags[0] = 1
        total = dd if k == 0 else (total - dd if A.subtract else total + dd)
        _arr(A.v_out, size).reshape(n)[region] = vnew[region]
        if A.vdot_out:
            _arr(A.vdot_out, size).reshape(n)[region] = rhs[region]
        if A.v_nan_count:
            _arr(A.v_nan_count, 1, np.int64)[0] += int(np.isnan(vnew[region]).sum())
    rhs = total * d.c_down
    pnew = p + rhs * d.dt
    _arr(d.p_out, size).reshape(n)[region] = pnew[region]
    if d.pdot_out:
        _arr(d.pdot_out, size).reshape(n)[region] = rhs[region]
    if d.p_nan_count:
        _arr(d.p_nan_count, 1, np.int64)[0] += int(np.isnan(pnew[region]).sum())
    return 0


FakeLib.fd_fused_pair_f64 = _pair


def _expr(self, desc, stream):
    """fd_fused_expr_f64 in numpy: run the postfix program on whole arrays, then
    reuse the fused step's epilogue through a one-term trick-free path."""
    self.launches += 1
    x = desc._obj if hasattr(desc, "_obj") else desc.contents
    d = x.base
    n = tuple(d.n)
    size = int(np.prod(n))
    flags = _arr(d.flags, 4, np.int64) if d.flags else None
    stack = []
    with np.errstate(all="ignore"):
        for pc in range(x.ncode):
            op, arg, imm = x.op[pc], x.arg[pc], x.imm[pc]
            if op == 0:
                stack.append(self._term(d.term[arg], n, flags))
            elif op == 1:
                stack.append(np.array(_arr(x.field[arg], size).reshape(n)))
            elif op == 2:
                stack.append(np.full(n, imm))
            elif op == 3:
                stack[-1] = -stack[-1]
            elif 4 <= op <= 8:
                b = stack.pop()
                a = stack.pop()
                stack.append([a + b, a - b, a * b, a / b, a ** b][op - 4] if False else
                             (a + b if op == 4 else a - b if op == 5 else a * b if op == 6 else a / b if op == 7 else a ** b))
            elif 9 <= op <= 13:
                a = stack[-1]
                stack[-1] = (a + imm if op == 9 else a - imm if op == 10 else a * imm if op == 11 else
                             a / imm if op == 12 else a ** imm)
            elif op == 14:
                stack[-1] = stack[-1] * stack[-1]
            elif op == 15:
                stack[-1] = 1 / stack[-1]
            elif op == 16:
                stack[-1] = np.sqrt(stack[-1])
    assert len(stack) == 1
    total = stack[0]
    region = tuple(slice(d.box_lo[a], d.box_hi[a]) for a in range(3)) if any(d.box_hi) else (slice(None),) * 3
    if d.dot_out:
        _arr(d.dot_out, size).reshape(n)[region] = total[region]
    rate = total
    for l in range(d.depth):
        new = _arr(d.prev[l], size).reshape(n) + rate * d.dt if d.prev[l] else np.array(rate, dtype=float)
        stored = np.full(n, np.nan)
        for dim in range(3):
            for side in range(2):
                fp = d.face[l][dim][side]
                if not fp:
                    continue
                fshape = [n[a] for a in range(3) if a != dim]
                face = _arr(fp, int(np.prod(fshape))).reshape(fshape)
                sel = [slice(None)] * 3
                sel[dim] = d.rim_lo[dim] if side == 0 else d.rim_hi[dim]
                tgt = stored[tuple(sel)]
                fill = np.isnan(tgt)
                tgt[fill] = face[fill]
        known = ~np.isnan(new)
        if flags is not None and np.any(~np.isnan(stored[region][known[region]])):
            flags[1] = 1
        new[~known] = stored[~known]
        _arr(d.out[l], size).reshape(n)[region] = new[region]
        if d.nan_count[l]:
            _arr(d.nan_count[l], 1, np.int64)[0] += int(np.isnan(new[region]).sum())
        rate = new
    return 0


FakeLib.fd_fused_expr_f64 = _expr


def _nested(self, inner_is_expr, inner, outer_is_expr, outer, scratch, scratch_elems, stream):
    """fd_fused_nested_f64: the scratch level goes where the descriptors say NULL."""
    import finite_difference._native as nat

    class _Ref:                      # what the fakes expect from a byref()
        def __init__(self, obj):
            self._obj = obj

    def copy(kind, address):
        cls = nat.Expr if kind else nat.Fused
        return cls.from_buffer_copy(cls.from_address(int(address)))

    i, o = copy(inner_is_expr, inner), copy(outer_is_expr, outer)
    ib, ob = (i.base if inner_is_expr else i), (o.base if outer_is_expr else o)
    assert ib.depth == 1 and not ib.prev[0] and not ib.out[0] and not ib.dot_out
    assert ib.n[0] * ib.n[1] * ib.n[2] <= scratch_elems
    ib.out[0] = scratch
    hits = 0
    for k in range(ob.nterms):
        if not ob.term[k].src:
            ob.term[k].src = scratch
            hits += 1
    if outer_is_expr:
        for k in range(o.nfields):
            if not o.field[k]:
                o.field[k] = scratch
                hits += 1
    assert hits
    rc = (self.fd_fused_expr_f64 if inner_is_expr else self.fd_fused_step_f64)(_Ref(i), stream)
    return rc or (self.fd_fused_expr_f64 if outer_is_expr else self.fd_fused_step_f64)(_Ref(o), stream)


FakeLib.fd_fused_nested_f64 = _nested


class FakeDevice:
    index = 0

    def __init__(self):
        self.dev = torch.device("cpu")
        self.copy_stream = None
        self.comm_stream = None
        self.rim_stream = None

    stream = None
    stream_ptr = 0

    def synchronize(self):
        pass

    def empty(self, n, dtype):
        return torch.empty(int(n), dtype=torch.float64)

    def zeros_i64(self, n):
        return torch.zeros(int(n), dtype=torch.int64)

    def upload(self, array, dtype, out=None):
        host = torch.from_numpy(np.ascontiguousarray(array, dtype=np.float64).reshape(-1).copy())
        if out is None:
            return host
        out.copy_(host)
        return out

    def download(self, tensor, shape=None):
        arr = tensor.numpy().copy()
        return arr.reshape(shape) if shape is not None else arr

    def pinned(self, n, dtype):
        return torch.empty(int(n), dtype=torch.float64)


@contextlib.contextmanager
def installed():
    """Route the package's native calls and device plumbing to the fakes."""
    import finite_difference._device as dev_mod
    import finite_difference._engine as eng_mod
    import finite_difference._native as nat

    fake = FakeLib()
    saved_fn, saved_get, saved_lib = nat.fn, dev_mod.Device.get, nat.lib
    fake_dev = FakeDevice()
    nat.fn = lambda name, dtype: getattr(fake, f"{name}_f64")
    nat.lib = fake
    dev_mod.Device.get = classmethod(lambda cls, index=None: fake_dev)
    eng_mod.Device = dev_mod.Device
    try:
        yield fake
    finally:
        nat.fn, nat.lib = saved_fn, saved_lib
        dev_mod.Device.get = saved_get
