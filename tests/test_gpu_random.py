"""Randomised consistency of the fused kernels (star / vector / scalar / pair,
interior-rim split) against the unfused operators, which execute the
reference's array operations one by one and are themselves checked against
the reference's golden outputs.  Results must be bit-identical."""

import contextlib
import io
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _build_and_run(fd, spec, fuse):
    rng = np.random.default_rng(spec["seed"])
    names = ["x", "y", "z"][:spec["ndim"]]
    axes = {"t": np.arange(spec["n_t"]) * spec["dt"]}
    for n, a, dx in zip(spec["shape"], names, spec["dx"]):
        axes[a] = np.arange(n) * dx
    os.environ["FD_B200_NO_FUSE"] = "0" if fuse else "1"
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model(axes, "t", periodic=tuple(a for a, p in zip(names, spec["periodic"]) if p),
                         dtype=spec["dtype"])
            u = fd.Field(m, "u", n_time_ders=spec["ders"])
            stencils = {a: fd.Stencil(list(range(-r, r + 1)), spec["order"]) for a, r in zip(names, spec["radius"])}
            u.set_IC_values(rng.random(spec["shape"]).astype(spec["dtype"]))
            if spec["ders"] == 2:
                u.dot.set_IC_values((rng.random(spec["shape"]) * 0.1).astype(spec["dtype"]))
            for a, p in zip(names, spec["periodic"]):
                if not p and a in spec["axes_used"]:   # a face without a NaN rim would be an override
                    u.set_BC(spec["bc"][0], a, "start")
                    u.set_BC(spec["bc"][1], a, "end")
            top = u
            for _ in range(spec["ders"]):
                top = top.dot
            while not m.finished:
                terms = [stencils[a].der(u.prev, a) for a in names if a in spec["axes_used"]]
                form = spec["form"]
                if form == 0:                      # c * (a + b + c)
                    acc = terms[0]
                    for tm in terms[1:]:
                        acc = acc + tm
                    expr = spec["c"] * acc
                elif form == 1:                    # c*a - c*b + ...
                    acc = terms[0] * spec["c"]
                    for i, tm in enumerate(terms[1:]):
                        acc = (acc - tm * spec["c"]) if i % 2 == 0 else (acc + tm * spec["c"])
                    expr = acc
                elif form == 2:                    # -(a + b) / s + 0.25
                    acc = terms[0]
                    for tm in terms[1:]:
                        acc = acc + tm
                    expr = -acc / 3.0 + 0.25
                else:                              # reversed order of the axes
                    acc = terms[-1]
                    for tm in reversed(terms[:-1]):
                        acc = acc + tm
                    expr = acc * spec["c"]
                top.assign_update(expr)
                f = top
                while f._parent is not None:
                    f = f._parent
                    f.time_integrate_update()
                m.increment_time()
            out = [np.array(u.data)]
            f = u
            while f.dot is not None:
                f = f.dot
                out.append(np.array(f.data))
        return out, dict(m._engine.stats)
    finally:
        os.environ.pop("FD_B200_NO_FUSE", None)


def _specs(n, seed):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        ndim = int(rng.integers(1, 4))
        zs = [int(rng.choice([8, 12, 16, 64, 130, 256, 384, 520]))]
        shape = [int(rng.integers(5, 40)) for _ in range(ndim - 1)] + zs
        if ndim == 2 and rng.random() < 0.5:
            shape = [int(rng.integers(300, 700)), int(rng.choice([256, 512, 640]))]
        radius = [int(rng.integers(1, 5))] * ndim if rng.random() < 0.7 else [int(rng.integers(1, 5)) for _ in range(ndim)]
        radius = [min(r, (n_ - 1) // 2) for r, n_ in zip(radius, shape)]
        if min(radius) < 1:
            continue
        names = ["x", "y", "z"][:ndim]
        used = [a for a in names if rng.random() < 0.85] or [names[-1]]
        out.append(dict(seed=1000 + i, ndim=ndim, shape=tuple(shape), radius=radius,
                        periodic=[bool(rng.integers(0, 2)) for _ in range(ndim)],
                        dx=[float(rng.choice([1.0, 1.0, 0.5, 0.3, 2.0])) for _ in range(ndim)],
                        order=int(rng.integers(1, 3)), ders=int(rng.integers(1, 3)), n_t=4,
                        dt=float(rng.choice([1e-3, 0.125])), c=float(rng.choice([0.1, -0.7, 2.0])),
                        form=int(rng.integers(0, 4)), axes_used=used,
                        bc=[str(rng.choice(["0", "1.5", "t", "t*2+1"])), str(rng.choice(["0", "-1", "t"]))],
                        dtype=np.float64 if rng.random() < 0.7 else np.float32))
    return out


@pytest.mark.parametrize("block", range(6))
def test_fused_equals_unfused_on_random_models(block):
    import finite_difference as fd
    specs = _specs(14, 4242 + block)
    assert len(specs) >= 8
    n_fused = 0
    for spec in specs:
        if spec["order"] >= 2 * min(spec["radius"]) + 1:
            spec["order"] = 1
        got, stats = _build_and_run(fd, spec, fuse=True)
        want, stats_u = _build_and_run(fd, spec, fuse=False)
        assert stats_u["fused"] == 0
        n_fused += stats["fused"] > 0
        for g, w in zip(got, want):
            assert np.array_equal(g, w, equal_nan=True), (spec, stats)
    assert n_fused >= len(specs) // 2


def _gray_scott(fd, shape, fuse, dtype=np.float64, steps=5):
    """Two coupled reaction-diffusion fields: products and powers of fields next
    to der terms (general point-wise expression kernel)."""
    os.environ["FD_B200_NO_FUSE"] = "0" if fuse else "1"
    try:
        rng = np.random.default_rng(7)
        names = ["x", "y", "z"][:len(shape)]
        axes = {"t": np.arange(steps + 1) * 0.05}
        for n, a in zip(shape, names):
            axes[a] = np.arange(n) * 0.5
        with contextlib.redirect_stdout(io.StringIO()):
            m = fd.Model(axes, "t", periodic=tuple(names), dtype=dtype)
            u = fd.Field(m, "u", n_time_ders=1)
            v = fd.Field(m, "v", n_time_ders=1)
            d2 = fd.Stencil([-1, 0, 1], 2)
            d1 = fd.Stencil([-1, 0, 1], 1)
            u.set_IC_values((0.5 + 0.5 * rng.random(shape)).astype(dtype))
            v.set_IC_values((0.25 * rng.random(shape)).astype(dtype))
            while not m.finished:
                lap_u = d2.der(u.prev, names[0])
                lap_v = d2.der(v.prev, names[0])
                for a in names[1:]:
                    lap_u = lap_u + d2.der(u.prev, a)
                    lap_v = lap_v + d2.der(v.prev, a)
                uvv = u.prev * v.prev ** 2
                u.dot.assign_update(0.16 * lap_u - uvv + 0.035 * (1 - u.prev) - u.prev * d1.der(u.prev, names[-1]) / 7)
                v.dot.assign_update(0.08 * lap_v + uvv - (0.035 + 0.065) * v.prev + (v.prev ** 0.5) * 1e-3)
                u.time_integrate_update()
                v.time_integrate_update()
                m.increment_time()
            return np.array(u.data), np.array(v.data), dict(m._engine.stats)
    finally:
        os.environ.pop("FD_B200_NO_FUSE", None)


@pytest.mark.parametrize("shape", [(300,), (65, 130), (20, 24, 70)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_expression_kernel_equals_unfused(shape, dtype):
    import finite_difference as fd
    u1, v1, st1 = _gray_scott(fd, shape, True, dtype)
    u0, v0, st0 = _gray_scott(fd, shape, False, dtype)
    assert st1.get("fused_expr", 0) == 10 and st1["unfused_statements"] == 0, st1
    assert st0["fused"] == 0
    assert np.array_equal(u1, u0, equal_nan=True) and np.array_equal(v1, v0, equal_nan=True)
