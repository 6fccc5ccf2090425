"""Generate ``tests/golden/*`` by running the UNMODIFIED reference.

TEST INFRASTRUCTURE ONLY.  Run in the build container (needs
``/root/reference``):  ``python oracle/make_golden.py``.

Outputs
  tests/golden/scenarios.npz     every array returned by oracle/scenarios.py::ALL
  tests/golden/scenarios.json    captured stdout of every scenario
  tests/golden/der_cases.npz     random Stencil.der cases (inputs and outputs)
  tests/golden/der_cases.json    their parameters
  tests/golden/weights.json      Stencil weights table (FD.py:938-951)
  tests/golden/misuse.json       exception type and message of every oracle/misuse_cases.py case
"""

from __future__ import annotations

import contextlib
import io
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_loader, scenarios  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

WEIGHT_CASES = [
    ([-1, 0, 1], 2), ([-1, 0, 1], 1), ([-2, -1, 0, 1, 2], 2), ([-2, -1, 0, 1, 2], 1),
    ([-3, -2, -1, 0, 1, 2, 3], 2), (list(range(-4, 5)), 2), (list(range(-4, 5)), 1),
    ([-0.5, 0.5], 1), ([-1.5, -0.5, 0.5, 1.5], 1), ([-1.5, -0.5, 0.5, 1.5], 2),
    ([0, 1], 1), ([-1, 0], 1), ([0, 1, 2], 2), ([-2, -1, 0], 1), ([0, 1, 2, 3], 3),
    ([-2, -1, 0, 1, 2], 4), ([-2.5, -1.5, -0.5, 0.5, 1.5, 2.5], 1),
]


def der_case_params(n_cases=120, seed=1234):
    """Random but reproducible ``der`` cases: kind x periodicity x 1-3 D."""
    rng = np.random.default_rng(seed)
    cases = []
    for i in range(n_cases):
        kind = ["cc", "ce", "ec", "ee"][i % 4]
        periodic = bool((i // 4) % 2)
        ndim = 1 + (i // 8) % 3
        half = kind in ("ce", "ec")
        npts = int(rng.integers(2, 6))
        start = int(rng.integers(-3, 1))
        pts = [start + j + (0.5 if half else 0.0) for j in range(npts)]
        if not half:
            pts = [int(p) for p in pts]
        order = int(rng.integers(1, min(npts, 3)))
        axis = int(rng.integers(0, ndim))
        shape = [int(rng.integers(7, 12)) for _ in range(ndim)]
        dx = [1.0, 0.5, 0.3, 2.0, 4][int(rng.integers(0, 5))]
        nan_frac = [0.0, 0.0, 0.1][int(rng.integers(0, 3))]
        cases.append(dict(kind=kind, periodic=periodic, ndim=ndim, points=pts, order=order,
                          axis=axis, shape=shape, dx=dx, nan_frac=nan_frac, seed=int(10_000 + i)))
    return cases


def run_der_case(fd, c):
    names = ["x", "y", "z"][:c["ndim"]]
    dx = c["dx"]
    axes = {"t": [0, 1]}
    for n, a in zip(c["shape"], names):
        axes[a] = (np.arange(n) * dx) if not isinstance(dx, int) else range(0, n * dx, dx)
    der_axis = names[c["axis"]]
    m = fd.Model(axes, "t", periodic=(der_axis,) if c["periodic"] else ())
    from_edge = c["kind"][0] == "e"
    to_edge = c["kind"][1] == "e"
    f = fd.Field(m, "f", edge_axes=(der_axis,) if from_edge else (), n_time_ders=1)
    with contextlib.redirect_stdout(io.StringIO()):
        st = fd.Stencil(list(c["points"]), c["order"], "edge" if from_edge else "cell",
                        "edge" if to_edge else "cell")
    rng = np.random.default_rng(c["seed"])
    slice_shape = f.data.shape[1:]
    vals = rng.standard_normal(slice_shape)
    if c["nan_frac"] > 0:
        vals[rng.random(slice_shape) < c["nan_frac"]] = np.nan
    f.data[0] = vals
    f.IC_set = True
    out = st.der(f.prev, der_axis)
    return vals, np.array(out.data), list(out.edge_axes), list(st.weights)


def main():
    fd = ref_loader.load()
    os.makedirs(GOLD, exist_ok=True)

    arrays, prints = {}, {}
    for name, fn in scenarios.ALL.items():
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            out = fn(fd)
        prints[name] = buf.getvalue()
        for k, v in out.items():
            arrays[f"{name}/{k}"] = np.array(v, dtype=np.float64)
        print(f"{name}: " + ", ".join(f"{k}{tuple(v.shape)}" for k, v in out.items()))
    np.savez_compressed(os.path.join(GOLD, "scenarios.npz"), **arrays)
    with open(os.path.join(GOLD, "scenarios.json"), "w") as fh:
        json.dump(prints, fh, indent=1, sort_keys=True)

    cases = der_case_params()
    packed = {}
    for i, c in enumerate(cases):
        vin, vout, edges, weights = run_der_case(fd, c)
        packed[f"in{i}"] = vin
        packed[f"out{i}"] = vout
        c["out_edge_axes"] = edges
        c["weights"] = weights
    np.savez_compressed(os.path.join(GOLD, "der_cases.npz"), **packed)
    with open(os.path.join(GOLD, "der_cases.json"), "w") as fh:
        json.dump(cases, fh, indent=0)

    table = []
    for pts, order in WEIGHT_CASES:
        half = any(p % 1 for p in pts)
        with contextlib.redirect_stdout(io.StringIO()) as buf:
            st = fd.Stencil(list(pts), order, "cell", "edge" if half else "cell")
        table.append(dict(points=pts, order=order, weights=list(st.weights), text=buf.getvalue()))
    with open(os.path.join(GOLD, "weights.json"), "w") as fh:
        json.dump(table, fh, indent=0)
    from oracle import misuse_cases
    misuse = {}
    for name, fn in misuse_cases.CASES.items():
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                fn(fd)
            misuse[name] = None
        except BaseException as exc:      # noqa: BLE001 - the type is what is being recorded
            misuse[name] = [type(exc).__name__, str(exc)]
    with open(os.path.join(GOLD, "misuse.json"), "w") as fh:
        json.dump(misuse, fh, indent=0, sort_keys=True)
    print(len(misuse), "misuse cases;", sum(v is None for v in misuse.values()), "did not raise in the reference")
    print("wrote", GOLD)


if __name__ == "__main__":
    main()
