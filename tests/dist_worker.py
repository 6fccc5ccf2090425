"""One rank of the slab-decomposition tests (spawned by test_dist_gloo.py on
the CPU with gloo + the numpy stand-in for the native library, and by
test_gpu_dist.py on GPUs with NCCL + the real library).

usage: python tests/dist_worker.py <backend> <scenario> [<scenario> ...]
Environment: RANK, WORLD_SIZE, MASTER_ADDR, MASTER_PORT (torch.distributed).
"""

import contextlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402



def _nccl_options():
    """NCCL kernels on a high-priority stream, so that a halo exchange issued
    while the interior kernel still has blocks to dispatch is not queued behind it."""
    import torch.distributed as dist
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = True
    return opts


def peer_halo_case(fd, world):
    """Slabs thick enough for the eager exchange: the boundary planes travel
    through peer memory (fd_halo_push / fd_halo_wait) unless FD_B200_NO_P2P=1;
    the gathered result is the single-process oracle's, bit for bit."""
    import io
    from oracle import fd_numpy as O
    from tests.test_gpu_parity import _acoustic, _heat
    direct = os.environ.get("FD_B200_NO_P2P", "0") != "1"
    shape, n_t = (8 * world, 24, 64), 12
    ic = np.random.default_rng(71).random(shape)
    for periodic in (True, False):
        m, T = _heat(fd, shape, n_t, periodic, 0.1, 1e-3, ic)
        want, _ = O.run_heat(shape, n_t, periodic, 0.1, 1e-3, ic)
        assert np.array_equal(T.data, want, equal_nan=True)
        assert (m._engine.stats.get("peer_halo_exchanges", 0) >= n_t - 2) == direct, m._engine.stats
    m, p, v = _acoustic(fd, shape, n_t, True, ic.astype(np.float32), np.float32)
    want, _ = O.run_acoustic(shape, n_t, 1.0, 0.25, 0.1, ic.astype(np.float32), dtype=np.float32)
    assert np.array_equal(p.data, want.astype(np.float64), equal_nan=True)
    assert (m._engine.stats.get("peer_halo_exchanges", 0) >= n_t - 2) == direct, m._engine.stats
    # slabs of unequal thickness (one plane more on the first ranks): other ghost offsets in the neighbours' rings
    shape = (8 * world + world - 1, 24, 64)
    ic = np.random.default_rng(72).random(shape)
    for periodic in (True, False):
        m, T = _heat(fd, shape, n_t, periodic, 0.1, 1e-3, ic)
        want, _ = O.run_heat(shape, n_t, periodic, 0.1, 1e-3, ic)
        assert np.array_equal(T.data, want, equal_nan=True)
        assert (m._engine.stats.get("peer_halo_exchanges", 0) >= n_t - 2) == direct, m._engine.stats


def cross_diffusion_case(fd, world):
    """oracle/scenarios.py:cross_diffusion on slabs: the result is the golden one
    and only the field with a single ghost reader took the peer-memory path."""
    from oracle import scenarios
    from tests.scenario_check import run_and_compare
    made = []
    real_model = fd.Model

    def spy(*a, **k):
        made.append(real_model(*a, **k))
        return made[-1]
    fd.Model = spy
    try:
        run_and_compare("cross_diffusion", fd, check_prints=False)
    finally:
        fd.Model = real_model
    eng = made[-1]._engine
    by_name = {st.field.name: st for st in eng.states.values()}
    if os.environ.get("FD_B200_NO_P2P", "0") != "1":
        assert by_name["c"].peer, "c is read by its own statement only: peer-memory path expected"
        assert not by_name["u"].peer, "u's ghosts are read by two statements: it must not be pushed"
    assert eng.stats["unfused_statements"] == 0, eng.stats


def nested_on_slabs_case(fd):
    """oracle/scenarios.py:nested_cross_axis on slabs: ders of computed expressions
    whose outer der stays clear of the decomposed axis are fused there too (scratch
    level with the fields' ghost layout), one of them with an inner der ALONG the
    decomposed axis; the golden bits."""
    from tests.scenario_check import run_and_compare
    made = []
    real_model = fd.Model

    def spy(*a, **k):
        made.append(real_model(*a, **k))
        return made[-1]
    fd.Model = spy
    try:
        run_and_compare("nested_cross_axis", fd, check_prints=False)
    finally:
        fd.Model = real_model
    stats = made[-1]._engine.stats
    assert stats["unfused_statements"] == 0 and stats.get("fused_nested") == 8, stats


def folded_axes_on_slabs_case(fd, world):
    """A 4-D heat model cut into slabs: the leading axes merge into dimension 0 of
    the kernels' view, the decomposed one (with its ghost planes) outermost; one
    fused launch per statement (three with the boundary planes split off for the
    overlapped exchange), the oracle's bits."""
    from tests.nd_cases import run_case
    for shape, periodic in (((4 * world, 3, 4, 8), True), ((4 * world + 1, 2, 6, 4), True)):
        got, want, stats = run_case(fd, shape, periodic, 4, n_t=6)
        assert np.array_equal(got, want, equal_nan=True), shape
        assert stats["fused"] == 5 and stats["unfused_statements"] == 0, stats
        assert stats.get("halo_exchanges", 0) >= 5, stats


def main():
    backend, names = sys.argv[1], sys.argv[2:]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)) % torch.cuda.device_count())
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()), pg_options=_nccl_options())
    else:
        dist.init_process_group("gloo")
    import finite_difference as fd
    from tests.scenario_check import run_and_compare

    ctx = contextlib.nullcontext()
    if backend == "gloo":
        from tests import fake_native
        ctx = fake_native.installed()
    with ctx:
        for name in names:
            if name == "__peer_halo__":
                peer_halo_case(fd, world)
                continue
            if name == "__cross_diffusion__":
                cross_diffusion_case(fd, world)
                continue
            if name == "__nested_on_slabs__":
                nested_on_slabs_case(fd)
                continue
            if name == "__folded_on_slabs__":
                folded_axes_on_slabs_case(fd, world)
                continue
            results = run_and_compare(name, fd, check_prints=False)
            assert results
    # every rank must have seen a decomposed model
    m = fd.Model({"t": [0, 1], "x": np.arange(4 * world) * 1.0}, "t")
    assert m._slab is not None and m._slab.world == world
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}/{world} ok: {len(names)} scenarios")


if __name__ == "__main__":
    main()
