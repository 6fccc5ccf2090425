"""Slab decomposition across the GPUs of one box (one process per GPU).

The reference is a single-process numpy program; nothing in it is distributed.
Here the slowest spatial axis (the first non-time axis of the model) is cut
into contiguous slabs, one per rank.  Every rank runs the *same* user script;
fields hold the rank's slab, ``set_IC``/``set_BC`` evaluate their expressions
on the slab's coordinates, and ``Field.data`` gathers the global array.

On the device a level is stored with ``H`` ghost planes on either side of the
owned planes (``H`` = largest stencil reach along the slab axis).  Before a
level is differentiated along the slab axis its ghost planes are refreshed
with one ``isend/irecv`` pair per neighbour (NCCL over NVLink on the GPU box,
gloo in the CPU tests); planes are contiguous in memory, so nothing is packed.
The kernels then see a non-periodic axis whose valid range is the owned planes
(``fd_fused_t.range_*``), which keeps the arithmetic -- and therefore the
results -- identical to the single-GPU run.
"""

from __future__ import annotations

import os
import sys

import numpy as np


class Slab:
    def __init__(self, dist, axis, n, periodic):
        self.dist = dist
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()
        self.axis = axis                    # axis name
        self.n = int(n)                     # global number of cells along it
        self.periodic = bool(periodic)
        base, extra = divmod(self.n, self.world)
        counts = [base + (1 if r < extra else 0) for r in range(self.world)]
        self.counts = counts
        self.lo = sum(counts[:self.rank])
        self.hi = self.lo + counts[self.rank]
        self.first = self.rank == 0
        self.last = self.rank == self.world - 1
        self.backend = dist.get_backend()

    def owned(self, extra_edge):
        """Planes this rank owns: its cells, plus the closing edge on the last
        rank for fields that live on the edges of a fixed slab axis."""
        return self.hi - self.lo + (1 if (extra_edge and self.last) else 0)

    def owned_of(self, rank, extra_edge):
        return self.counts[rank] + (1 if (extra_edge and rank == self.world - 1) else 0)

    @property
    def prev_rank(self):
        if self.first:
            return self.world - 1 if self.periodic else None
        return self.rank - 1

    @property
    def next_rank(self):
        if self.last:
            return 0 if self.periodic else None
        return self.rank + 1

    # -- collectives on small host/device words ------------------------------
    def all_reduce(self, tensor, op):
        ops = {"max": self.dist.ReduceOp.MAX, "sum": self.dist.ReduceOp.SUM}
        self.dist.all_reduce(tensor, op=ops[op])
        return tensor

    def gather_arrays(self, local, axis):
        """Concatenate every rank's array along ``axis`` (host side; used by
        ``Field.data``)."""
        parts = [None] * self.world
        self.dist.all_gather_object(parts, np.ascontiguousarray(local))
        return np.concatenate(parts, axis=axis)

    def exchange(self, tensor, plane, owned, depth):
        """Refresh the ``depth`` ghost planes on both sides of a level laid out as
        [depth ghosts | owned planes | depth ghosts], each plane ``plane`` elements."""
        if depth == 0:
            return
        d, P = self.dist, plane
        lo_ghost = tensor[0:depth * P]
        lo_own = tensor[depth * P:2 * depth * P]
        hi_own = tensor[owned * P:(owned + depth) * P]
        hi_ghost = tensor[(depth + owned) * P:(2 * depth + owned) * P]
        ops = []
        prev, nxt = self.prev_rank, self.next_rank
        # sends first (lo planes to the previous rank, hi planes to the next),
        # receives in the order the peers send: when both neighbours are the same
        # rank (world == 2, periodic) messages are matched by order alone
        if prev is not None:
            ops.append(d.P2POp(d.isend, lo_own, prev))
        if nxt is not None:
            ops.append(d.P2POp(d.isend, hi_own, nxt))
        if nxt is not None:
            ops.append(d.P2POp(d.irecv, hi_ghost, nxt))
        if prev is not None:
            ops.append(d.P2POp(d.irecv, lo_ghost, prev))
        if ops:
            for req in d.batch_isend_irecv(ops):
                req.wait()


def make_slab(model):
    """Slab description for ``model`` if this process is one of several ranks."""
    if os.environ.get("FD_B200_DIST", "1") == "0":
        return None
    if "torch" not in sys.modules and os.environ.get("WORLD_SIZE", "1") in ("", "1"):
        return None
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        return None
    space = [a for a in model._names if a != model.time_axis]
    if not space:
        return None
    axis = space[0]
    n = model._lengths[model._names.index(axis)]
    if n < 2 * dist.get_world_size():
        raise Exception(f"axis {axis} has {n} points, too few to cut into {dist.get_world_size()} slabs")
    return Slab(dist, axis, n, axis in model.periodic)
