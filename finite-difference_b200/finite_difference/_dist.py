"""Slab decomposition across the GPUs of one box (one process per GPU).

The reference is a single-process numpy program; nothing in it is distributed.
Here the slowest spatial axis (the first non-time axis of the model) is cut
into contiguous slabs, one per rank.  Every rank runs the *same* user script;
fields hold the rank's slab, ``set_IC``/``set_BC`` evaluate their expressions
on the slab's coordinates, and ``Field.data`` gathers the global array.

On the device a level is stored with ``H`` ghost planes on either side of the
owned planes (``H`` = largest stencil reach along the slab axis).  Before a
level is differentiated along the slab axis its ghost planes are refreshed:
on the GPU box the neighbours write them straight into this rank's memory
over NVLink (``PeerLinks`` below, csrc/fd_halo.cu); where that is not possible
-- and in the CPU tests (gloo) -- with one ``isend/irecv`` pair per neighbour.
Planes are contiguous in memory, so nothing is packed.
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
        ops = {"max": self.dist.ReduceOp.MAX, "sum": self.dist.ReduceOp.SUM, "min": self.dist.ReduceOp.MIN}
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


class RingPool:
    """Rings whose neighbours hold mappings of them, between two models.

    A ring that a peer has mapped must not go back to the allocator while the
    peer can still store into it, and mapping it again costs two object
    collectives plus the IPC opens.  A closed model therefore leaves such rings
    here (``Engine.close``), keyed by their geometry, and the next model of the
    same geometry takes them over with flag words and mappings intact
    (``Engine.ring_pool_take``; whether every rank found one is agreed with one
    all-reduce before the mappings are trusted).  Entries live until the process
    ends."""

    _free = {}

    @classmethod
    def take(cls, key):
        got = cls._free.get(key)
        return got.pop() if got else None

    @classmethod
    def give(cls, key, bundle):
        cls._free.setdefault(key, []).append(bundle)


class PeerLinks:
    """Peer-memory halo exchange of one field's ring (NVLink / NVSwitch box).

    Every rank maps the level buffers and a few flag words of its two
    neighbours (CUDA IPC through the C ABI: ``fd_ipc_export`` / ``fd_ipc_open``).
    ``push`` then writes this rank's boundary planes straight into the
    neighbours' ghost planes and publishes the level number in their flag
    words; ``wait`` holds the stream until the neighbours' planes have arrived
    here.  No communication library on the data path (NCCL's host-side cost per
    isend/irecv batch is ~60 us, the push+wait pair takes ~20 us for 2 MiB
    planes, profiles/p2p_probe.py).

    Set-up is collective (handles travel with ``all_gather_object``): every rank
    must create the links of its fields in the same order.

    The flag words carry ``base + level``: a level number only grows within a
    model, and ``retire`` moves ``base`` past everything the closed model can
    have published, so the next model that takes the ring over (``RingPool``)
    starts above every stale value without the words being reset.
    """

    TIMEOUT_NS = 20_000_000_000
    _opened = {}          # (rank, handle) -> mapped base address, per process

    def __init__(self, slab, nat, ring, flags, plane_bytes, owned0, H, base=0):
        import ctypes as C
        self.nat, self.H = nat, H
        self.base = int(base)
        self.nbytes = H * plane_bytes
        self.flags = flags                     # int64[len(ring) * 2]: [slot][0: low ghost fresh, 1: high ghost fresh]
        lib = nat.lib

        def export(t):
            h = C.create_string_buffer(nat.FD_IPC_HANDLE_BYTES)
            off = C.c_longlong()
            nat.check(lib.fd_ipc_export(t.data_ptr(), h, C.byref(off)))
            return bytes(h.raw), off.value

        ok, mine = True, None
        try:
            if self.nbytes % 16 or plane_bytes % 16:
                raise nat.NativeError("planes are not multiples of 16 bytes")
            mine = {"ring": [export(t) for t in ring], "flags": export(flags), "owned0": owned0}
        except nat.NativeError:
            ok = False
        everyone = [None] * slab.world
        slab.dist.all_gather_object(everyone, mine)
        ok = all(e is not None for e in everyone)
        prev, nxt = slab.prev_rank, slab.next_rank
        self.src_lo, self.src_hi, self.dst_lo, self.dst_hi = [], [], [], []
        self.flag_lo, self.flag_hi, self.my_lo, self.my_hi = [], [], [], []
        if ok:
            try:
                for k, t in enumerate(ring):
                    base = t.data_ptr()
                    # my low owned planes -> prev's high ghost planes; my high owned planes -> next's low ghost planes
                    self.src_lo.append(base + H * plane_bytes if prev is not None else None)
                    self.src_hi.append(base + owned0 * plane_bytes if nxt is not None else None)
                    self.dst_lo.append(self._map(prev, everyone, ("ring", k)) + (H + everyone[prev]["owned0"]) * plane_bytes
                                       if prev is not None else None)
                    self.dst_hi.append(self._map(nxt, everyone, ("ring", k)) if nxt is not None else None)
                    self.flag_lo.append(self._map(prev, everyone, ("flags",)) + (2 * k + 1) * 8 if prev is not None else None)
                    self.flag_hi.append(self._map(nxt, everyone, ("flags",)) + (2 * k) * 8 if nxt is not None else None)
                    self.my_lo.append(flags.data_ptr() + (2 * k) * 8 if prev is not None else None)
                    self.my_hi.append(flags.data_ptr() + (2 * k + 1) * 8 if nxt is not None else None)
            except nat.NativeError:
                ok = False
        # every rank must take the same path: one failed mapping sends all of them back to isend/irecv
        verdicts = [None] * slab.world
        slab.dist.all_gather_object(verdicts, ok)
        self.ok = all(verdicts)

    def _map(self, rank, everyone, key):
        import ctypes as C
        entry = everyone[rank][key[0]]
        handle, off = entry[key[1]] if len(key) > 1 else entry
        got = PeerLinks._opened.get((rank, handle))
        if got is None:
            if not PeerLinks._opened:
                import atexit
                atexit.register(PeerLinks.close_all, self.nat)
            base = C.c_void_p()
            self.nat.check(self.nat.lib.fd_ipc_open(handle, C.byref(base)))
            got = PeerLinks._opened[(rank, handle)] = base.value
        return got + off

    def retire(self, n_levels):
        """The model that used these links is closed (same call on every rank)."""
        self.base += int(n_levels) + 2

    def push(self, k, value, stream):
        self.nat.check(self.nat.lib.fd_halo_push(self.src_lo[k], self.dst_lo[k], self.src_hi[k], self.dst_hi[k], self.nbytes,
                                                 self.flag_lo[k], self.flag_hi[k], self.base + value, stream))

    def wait(self, k, value, err_flags, stream):
        self.nat.check(self.nat.lib.fd_halo_wait(self.my_lo[k], self.my_hi[k], self.base + value, self.TIMEOUT_NS,
                                                 err_flags, stream))

    def fill(self, link, k, value):
        """Describe the exchange of ring slot ``k`` at level number ``value`` in a
        ``fd_halo_link_t`` (fd_halo_exchange / fd_slab_step)."""
        link.src_lo, link.peer_dst_lo = self.src_lo[k], self.dst_lo[k]
        link.src_hi, link.peer_dst_hi = self.src_hi[k], self.dst_hi[k]
        link.nbytes = self.nbytes
        link.peer_flag_lo, link.peer_flag_hi = self.flag_lo[k], self.flag_hi[k]
        link.my_flag_lo, link.my_flag_hi = self.my_lo[k], self.my_hi[k]
        link.value = self.base + value

    @classmethod
    def close_all(cls, nat):
        """Unmap every neighbour buffer this process opened (process exit)."""
        opened, cls._opened = cls._opened, {}
        for base in opened.values():
            try:
                nat.lib.fd_ipc_close(base)
            except Exception:
                pass
