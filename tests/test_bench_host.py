"""Host-side pieces of bench.py that need no GPU: the clock helper's windowing,
the initial-value filler, and the reference arm's step functions (the unmodified
FD.py where it is present, the numpy port otherwise) on small grids."""

import numpy as np
import pytest

import bench
from oracle import ref_loader


def test_clock_summary_selects_the_samples_of_the_timed_region():
    helper = bench.ClockHelper([])                 # no process: rows filled in by hand
    helper.source = "nvml"
    helper.rows = [[10.0 + 0.001 * i, g, 1965 - 100 * (i > 50), 1965, 0x4 if i == 30 else 0]
                   for i in range(100) for g in (0, 1)]
    inside = helper.summary((10.010, 10.040))
    assert inside["samples"] == 2 * 31 and inside["sm_mhz"] == 1965.0 and inside["reasons"] == ["sw_power_cap"]
    assert set(inside["per_gpu"]) == {"0", "1"}
    later = helper.summary((10.060, 10.080))
    assert later["sm_mhz"] == 1865.0 and later["reasons"] == []
    # a region shorter than one sampling round: the nearest sample of every GPU
    tiny = helper.summary((10.0204, 10.0206))
    assert tiny["samples"] == 2
    empty = bench.ClockHelper([])
    empty.rows = []
    assert empty.summary((0.0, 1.0))["samples"] == 0


def test_fill_ic_small_and_large_levels():
    small = np.empty((4, 5, 6))
    bench.fill_ic(small, 3)
    assert np.array_equal(small, np.random.default_rng(3).random((4, 5, 6)))
    big = np.empty((3, 1 << 14, 1 << 12), dtype=np.float32)       # above 2^27 points: one plane repeated
    bench.fill_ic(big, 1)
    assert np.array_equal(big[0], big[2]) and 0.0 <= big.min() and big.max() <= 1.0     # (fp32 rounding can reach 1.0)


@pytest.mark.parametrize("workload,shape", [("c3", (12, 10, 8)), ("c2", (20, 24)), ("c4", (24, 20))])
def test_reference_arm_port_and_live_reference_agree(workload, shape):
    """Both step functions of the reference arm advance the same state the same
    way (chained 4-level models against one long numpy run)."""
    if not ref_loader.available():
        pytest.skip("reference source not present on this machine")
    rng = np.random.default_rng(8)
    state = (rng.random(shape), np.zeros(shape) if workload == "c4" else None)
    kind, live = bench.reference_step_fn(workload, shape)
    assert kind == "reference"
    saved = ref_loader.available
    ref_loader.available = lambda: False
    try:
        kind, port = bench.reference_step_fn(workload, shape)
    finally:
        ref_loader.available = saved
    assert kind == "port"
    _, a = live(7, state)
    _, b = port(7, state)
    assert np.array_equal(a[0], b[0], equal_nan=True)
    if workload == "c4":
        assert np.array_equal(a[1], b[1], equal_nan=True)
