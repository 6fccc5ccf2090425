"""Slab decomposition on the CPU: world_size 2 and 3 with the gloo backend and
the numpy stand-in for the native library.  Every rank runs the unchanged
scenario scripts; the gathered Field.data must equal the reference's golden
output bit for bit, exactly like the single-process run."""

import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCENARIOS = ["heat3d_periodic", "heat3d_fixed", "heat2d_fixed", "heat2d_time_middle", "wave2d", "wave2d_5pt",
             "wave2d_fixed_nan_spread", "acoustic3d", "acoustic3d_fixed", "flux_1d", "variable_coefficient",
             "burgers_1d", "readme_1d", "acoustic3d_wide", "acoustic2d", "heat4d_periodic", "heat1d_odd_dx",
             "cross_diffusion", "nested_cross_axis"]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(world, backend, names, extra_env=None, timeout=180, attempts=3):
    """Run the worker on `world` ranks.  Rendezvous hiccups (a port grabbed
    between probing and binding, a slow peer) are retried on a fresh port; a
    genuine failure fails every attempt."""
    for attempt in range(attempts):
        try:
            return _launch_once(world, backend, names, extra_env, timeout)
        except AssertionError as exc:
            if attempt == attempts - 1:
                raise
            print(f"attempt {attempt + 1} failed, retrying on a fresh port:\n{str(exc)[-800:]}")


def _launch_once(world, backend, names, extra_env, timeout):
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), **(extra_env or {}))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), backend, *names],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise AssertionError("rendezvous or run timed out")
        outs.append(out)
    if any(p.returncode != 0 for p in procs):
        raise AssertionError("\n".join(f"--- rank {r} (exit {p.returncode}) ---\n{o[-2500:]}"
                                       for r, (p, o) in enumerate(zip(procs, outs))))
    for rank, (p, out) in enumerate(zip(procs, outs)):
        assert f"rank {rank}/{world} ok" in out


@pytest.mark.parametrize("world", [2, 3])
def test_slab_decomposition_gloo(world):
    names = [n for n in SCENARIOS if not (world == 3 and n in ("acoustic3d_fixed", "heat4d_periodic"))]
    launch(world, "gloo", names)


@pytest.mark.parametrize("world", [2, 3])
def test_nested_ders_are_fused_on_slabs(world):
    launch(world, "gloo", ["__nested_on_slabs__"])


@pytest.mark.parametrize("world", [2, 3])
def test_more_than_three_axes_are_fused_on_slabs(world):
    launch(world, "gloo", ["__folded_on_slabs__"])


def test_slab_decomposition_gloo_unfused():
    launch(2, "gloo", ["heat3d_fixed", "wave2d", "acoustic3d_fixed", "flux_1d", "cross_diffusion"],
           {"FD_B200_NO_FUSE": "1"})
