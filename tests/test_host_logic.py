"""Host-side logic of the package on the CPU-only container: the same
scenarios as the GPU parity tests, with libfd_b200.so and the device replaced
by the numpy stand-ins of tests/fake_native.py (test infrastructure)."""

import pytest

import finite_difference as fd
from oracle import scenarios
from tests import fake_native
from tests.scenario_check import run_and_compare


@pytest.mark.parametrize("name", sorted(scenarios.ALL))
def test_scenario_fake_native(name):
    with fake_native.installed() as fake:
        run_and_compare(name, fd)
        assert fake.launches > 0


@pytest.mark.parametrize("name", ["heat2d_fixed", "wave2d", "acoustic3d_fixed", "flux_1d"])
def test_scenario_unfused_only(name, monkeypatch):
    monkeypatch.setenv("FD_B200_NO_FUSE", "1")
    with fake_native.installed():
        run_and_compare(name, fd)
