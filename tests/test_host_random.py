"""Randomised check of the host planner on the CPU: the launch descriptors the
engine builds for random models (fused path) must give the same arrays as its
operator-by-operator path, with the numpy stand-in of tests/fake_native.py
executing both (test infrastructure; fp64 only).  The GPU twin of this test is
tests/test_gpu_random.py."""

import numpy as np
import pytest

from tests import fake_native
from tests.test_gpu_random import _build_and_run, _specs


@pytest.mark.parametrize("block", range(3))
def test_fused_descriptors_equal_unfused_operators(block):
    n = fused = 0
    with fake_native.installed():
        import finite_difference as fd
        for spec in _specs(14, 9000 + block):
            if spec["dtype"] != np.float64:
                continue
            if max(spec["shape"]) > 300:            # keep the numpy stand-in quick
                spec["shape"] = tuple(min(s, 96) for s in spec["shape"])
                spec["radius"] = [min(r, (s - 1) // 2) for r, s in zip(spec["radius"], spec["shape"])]
            if spec["order"] >= 2 * min(spec["radius"]) + 1:
                spec["order"] = 1
            got, stats = _build_and_run(fd, spec, True)
            want, stats_u = _build_and_run(fd, spec, False)
            assert stats_u["fused"] == 0
            n += 1
            fused += stats["fused"] > 0
            for g, w in zip(got, want):
                assert np.array_equal(g, w, equal_nan=True), (spec, stats)
    assert n >= 5 and fused >= n // 2
