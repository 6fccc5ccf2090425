"""Shared by the CPU (fake native) and GPU parity tests: run one scenario of
oracle/scenarios.py through the B200 package and compare every returned array
bit for bit (NaN pattern included) with the reference's golden output."""

import contextlib
import io
import json
import os

import numpy as np

from oracle import scenarios

GOLD = os.path.join(os.path.dirname(__file__), "golden")
_cache = {}


def golden():
    if not _cache:
        _cache["arrays"] = np.load(os.path.join(GOLD, "scenarios.npz"))
        _cache["prints"] = json.load(open(os.path.join(GOLD, "scenarios.json")))
    return _cache["arrays"], _cache["prints"]


def run_and_compare(name, fd, check_prints=True):
    arrays, prints = golden()
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        out = scenarios.ALL[name](fd)
        results = {k: np.array(v) for k, v in out.items()}
    for key, got in results.items():
        want = arrays[f"{name}/{key}"]
        assert got.shape == want.shape, (name, key, got.shape, want.shape)
        same = np.array_equal(got, want, equal_nan=True)
        if not same:
            bad = ~((got == want) | (np.isnan(got) & np.isnan(want)))
            first = tuple(int(i[0]) for i in np.nonzero(bad))
            raise AssertionError(f"{name}/{key}: {int(bad.sum())} of {bad.size} values differ; first at {first}: "
                                 f"got {got[first]!r}, want {want[first]!r}")
    if check_prints:
        assert buf.getvalue() == prints[name], (name, buf.getvalue(), prints[name])
    return results
