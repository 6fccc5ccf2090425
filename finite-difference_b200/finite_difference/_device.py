"""Device plumbing: memory, streams, events and flag words, via PyTorch.

PyTorch is used here only as a CUDA allocator / stream / event provider (and,
for multi-GPU runs, for its NCCL process group).  All arithmetic happens in
``libfd_b200.so``.  Nothing in this module computes on the CPU; if no CUDA
device is available every entry point raises.
"""

from __future__ import annotations

import os

import numpy as np

from . import _native as nat

_torch = None


def torch():
    global _torch
    if _torch is None:
        import torch as _t
        _torch = _t
    return _torch


class NoDeviceError(RuntimeError):
    pass


def _torch_dtype(dtype):
    t = torch()
    return t.float32 if np.dtype(dtype) == np.float32 else t.float64


class Device:
    """One CUDA device: a compute stream, a copy (side) stream and the sticky
    condition flags shared by every launch of a model."""

    _instances = {}

    @classmethod
    def get(cls, index=None):
        t = torch()
        if not t.cuda.is_available():
            raise NoDeviceError(
                "finite_difference (B200 build) needs a CUDA device to advance a model; "
                "there is no CPU fallback")
        if index is None:
            index = int(os.environ.get("LOCAL_RANK", "0")) % t.cuda.device_count() \
                if "LOCAL_RANK" in os.environ else t.cuda.current_device()
        if index not in cls._instances:
            cls._instances[index] = cls(index)
        return cls._instances[index]

    def __init__(self, index):
        t = torch()
        self.index = index
        self.dev = t.device("cuda", index)
        t.cuda.set_device(self.dev)
        # kernels run on torch's current stream so that torch.cuda.Event timing
        # and NCCL ordering see them; copies to the host run on a side stream
        self.copy_stream = t.cuda.Stream(device=self.dev)
        # halo exchanges of a slab-decomposed model run here, ahead of the interior kernel
        self.comm_stream = t.cuda.Stream(device=self.dev, priority=-1)
        # the thin rim slabs of a fixed-boundary statement run here, next to the interior box
        self.rim_stream = t.cuda.Stream(device=self.dev, priority=-1)

    # -- streams ---------------------------------------------------------
    @property
    def stream(self):
        return torch().cuda.current_stream(self.dev)

    @property
    def stream_ptr(self):
        return self.stream.cuda_stream

    def synchronize(self):
        torch().cuda.synchronize(self.dev)

    # -- memory ----------------------------------------------------------
    def empty(self, n, dtype):
        return torch().empty(int(n), dtype=_torch_dtype(dtype), device=self.dev)

    def zeros_i64(self, n):
        t = torch()
        return t.zeros(int(n), dtype=t.int64, device=self.dev)

    def ones_i64(self, n):
        t = torch()
        return t.ones(int(n), dtype=t.int64, device=self.dev)

    def upload(self, array, dtype, out=None):
        """Host ndarray -> device tensor (flattened, cast to ``dtype``)."""
        t = torch()
        host = np.ascontiguousarray(array, dtype=dtype).reshape(-1)
        src = t.from_numpy(host)
        if out is None:
            out = self.empty(host.size, dtype)
        out.copy_(src, non_blocking=False)
        return out

    def download(self, tensor, shape=None):
        arr = tensor.detach().cpu().numpy()
        return arr.reshape(shape) if shape is not None else arr

    # -- pinned host memory for snapshots --------------------------------------
    ARENA_BYTES = 8 << 20

    def pinned(self, n, dtype):
        """Flat pinned host tensor of ``n`` elements.  Small requests are carved
        out of 8 MiB arenas (one cudaHostAlloc per arena, not per time level);
        the arena lives as long as any view of it does."""
        t = torch()
        td = _torch_dtype(dtype)
        nbytes = int(n) * np.dtype(dtype).itemsize
        if nbytes >= self.ARENA_BYTES // 4:
            return self._host_alloc(int(n), td)
        arena = getattr(self, "_arena", None)
        if arena is None or arena[2] != td or arena[1] + int(n) > arena[0].numel():
            arena = [self._host_alloc(self.ARENA_BYTES // np.dtype(dtype).itemsize, td), 0, td]
            self._arena = arena
        view = arena[0][arena[1]:arena[1] + int(n)]
        arena[1] += (int(n) + 31) // 32 * 32          # keep views 256-byte aligned
        return view

    @staticmethod
    def _host_alloc(n, td):
        t = torch()
        try:
            return t.empty(n, dtype=td, pin_memory=True)
        except RuntimeError:                             # pinning refused: pageable still works
            return t.empty(n, dtype=td)

    def event(self):
        return torch().cuda.Event()


def ptr(tensor):
    return 0 if tensor is None else tensor.data_ptr()


def launch_count():
    return int(nat.lib.fd_launch_count())
