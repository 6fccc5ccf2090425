"""Slab decomposition on real GPUs (NCCL halo exchange over NVLink): needs at
least two devices, so it is skipped on the single-GPU box and run with
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_dist.py -m gpu`."""

import pytest

from tests.test_dist_gloo import SCENARIOS, launch

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4])
def test_slab_decomposition_nccl(world):
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    names = [n for n in SCENARIOS if not (world == 4 and n in ("acoustic3d_fixed", "heat4d_periodic", "heat3d_periodic",
                                                               "heat3d_fixed", "acoustic3d"))]
    launch(world, "nccl", names)


def test_slab_decomposition_nccl_unfused():
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    launch(2, "nccl", ["heat3d_fixed", "wave2d", "acoustic3d_fixed", "flux_1d", "cross_diffusion"],
           {"FD_B200_NO_FUSE": "1"})


def test_more_than_three_axes_are_fused_on_slabs_nccl():
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    launch(2, "nccl", ["__folded_on_slabs__"])


def test_nested_ders_are_fused_on_slabs_nccl():
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    launch(2, "nccl", ["__nested_on_slabs__"])


@pytest.mark.parametrize("world", [2, 4])
def test_two_statements_read_the_same_ghost_planes(world):
    """`u`'s current level is differentiated along the slab axis by its own
    statement and by `c`'s: a neighbour's push of level t+2 could land in the
    slot whose level-t ghosts the second statement has yet to read, so `u` must
    travel by isend/irecv (a rendezvous) while `c` may use peer memory.  The
    worker checks the engine's choice and the bits."""
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    launch(world, "nccl", ["__cross_diffusion__"])


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("direct", [True, False])
def test_peer_memory_halo_exchange(world, direct):
    """Boundary planes written straight into the neighbours' ghost planes
    (CUDA-IPC mapped rings, fd_halo_push / fd_halo_wait) and, with
    FD_B200_NO_P2P=1, through NCCL isend/irecv: same bits as one process."""
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    launch(world, "nccl", ["__peer_halo__"], {} if direct else {"FD_B200_NO_P2P": "1"})
