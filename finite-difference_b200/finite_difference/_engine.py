"""GPU execution of the statements recorded during one time step.

For every ``assign_update`` / ``time_integrate_update`` pair (or triple, for
``n_time_ders=2``) the engine first tries the **fused statement kernel**
(``fd_fused_step``: der terms + scalar arithmetic + Euler chain + boundary
merge in one pass over HBM, FD.py:990-1122 + 794-869 + 478-576).  Anything the
pattern matcher does not recognise runs through the **unfused operators**
(``fd_der``, ``fd_ewise``, ``fd_merge_store``), one launch per reference array
operation.  Both produce the reference's values bit for bit.

Device layout: every field owns a small ring of time levels (2 for fields
that are integrated in time, 1 for everything else); a level is the field's
spatial slice in C order, the last spatial axis contiguous.  Boundary faces
set with ``set_BC`` live next to the ring (one copy if they do not depend on
time, ``n_t`` copies otherwise) and are merged in-kernel.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _native as nat
from ._device import Device, ptr
from ._graph import Binary, Const, Der, Leaf, Node, Unary

INSTABILITY_MSG = "Instability has led to infinite values, considering decreasing timestep"  # FD.py:1113
OVERRIDE_MSG = "attempting to override values, this may be due to too many boundary conditions."  # FD.py:572
LOSING_MSG = "Warning: fields seem to be losing data over time\n\
This is possibly due to insufficent boundary conditions\n"  # FD.py:536-537

_FS_OPS = {"add": nat.OP_ADD_FS, "sub": nat.OP_SUB_FS, "mul": nat.OP_MUL_FS, "truediv": nat.OP_DIV_FS}
_FF_OPS = {"add": nat.OP_ADD_FF, "sub": nat.OP_SUB_FF, "mul": nat.OP_MUL_FF, "truediv": nat.OP_DIV_FF,
           "pow": nat.OP_POW_FF}
_POST_OPS = {"add": nat.POST_ADD, "sub": nat.POST_SUB, "mul": nat.POST_MUL, "truediv": nat.POST_DIV}


class FieldState:
    """Device-side state of one field."""

    def __init__(self, engine, field):
        from .field import ConstantField
        self.field = field
        self.engine = engine
        self.size = int(np.prod(field._slice_shape)) if field._slice_shape else 1   # owned elements
        self.constant = isinstance(field, ConstantField)
        self.integrated = field.dot is not None
        depth = 2 if self.integrated else 1
        # slab decomposition: [H ghost planes | owned planes | H ghost planes]
        self.H = engine.H
        shape = tuple(field._slice_shape)
        self.owned0 = shape[0] if shape else 1
        self.plane = self.size // self.owned0 if self.owned0 else 0
        self.ext_shape = ((self.owned0 + 2 * self.H,) + shape[1:]) if self.H else shape
        self.ext_size = self.size + 2 * self.H * self.plane
        # (several GPUs: a ring whose peer mappings survive from an earlier model of the same geometry)
        self.pooled = engine.ring_pool_take(self, depth)
        self.ring = (self.pooled["ring"] if self.pooled is not None
                     else [engine.dev.empty(self.ext_size, engine.dtype) for _ in range(depth)])
        self.held = [-1] * depth          # time index each slot holds
        self.halo_ok = [False] * depth    # ghost planes of the slot are current
        self.halo_event = [None] * depth  # ... once this event (comm stream) has completed
        self.wants_ghosts = False         # some statement differentiates this field along the slab axis
        self.peer = None                  # PeerLinks once mapped, False if peer memory is unavailable
        self.saved = set()                # time indices already in the host history
        self.faces = None                 # {(dim, side): (tensor, per_time_elems or 0)}
        self.counts = engine.dev.zeros_i64(engine.n_t) if self.integrated else None
        self.counts_checked = engine.model.time_step
        s = len(field._slice_shape)
        if s <= 3:
            self.dims3 = (1,) * (3 - s) + tuple(self.ext_shape)
        else:
            # more than three spatial axes: the leading ones merge into dimension 0
            # of the kernels' view (fd_term_t.fold_div / fold_mod); on a slab the
            # outermost of them carries the ghost planes
            self.dims3 = (int(np.prod(self.ext_shape[:-2])),) + tuple(self.ext_shape[-2:])
        if not self.constant:
            self._discover_array_faces()
        if self.constant:
            self.load_level(0)
        elif self.integrated:
            self.load_level(engine.model.time_step)

    def slot_index(self, t):
        return t % len(self.ring)

    def _discover_array_faces(self):
        """Boundary values written straight into ``Field.data`` (set_BC only takes
        strings, so array-valued boundaries are written into the future time levels
        by hand, exactly like array initial conditions): any boundary face of a
        fixed axis that holds known values at a future level is treated like a
        set_BC face.  Known future values anywhere else cannot be honoured."""
        f, model = self.field, self.engine.model
        dense = f._hist.dense
        t_from = model.time_step + 1
        if dense is None or t_from >= self.engine.n_t:
            return
        t_pos = model._names.index(model.time_axis)
        future = np.moveaxis(dense, t_pos, 0)[t_from:]
        known = ~np.isnan(future)
        if not known.any():
            return
        space = [a for a in model._names if a != model.time_axis]
        for k, axis in enumerate(space):
            if axis in model.periodic:
                continue
            for idx in (0, -1):
                sel = [slice(None)] * future.ndim
                sel[1 + k] = idx
                if known[tuple(sel)].any() and (axis, idx) not in f._bc_locations:
                    f._bc_locations.append((axis, idx))
        for axis, idx in f._bc_locations:
            sel = [slice(None)] * future.ndim
            sel[1 + space.index(axis)] = idx
            known[tuple(sel)] = False
        if known.any():
            raise NotImplementedError(
                f"field {f.name}: values were written into future time levels of Field.data away from the "
                "boundary faces; the B200 engine only pre-populates boundary faces")

    def load_level(self, t):
        """Host history -> device (initial condition, constant coefficient)."""
        eng = self.engine
        k = self.slot_index(t)
        uniform = self.field._hist.uniform_level(t)
        if uniform is not None:             # e.g. set_IC("0"): filled on the device
            nat.check(eng._fn["fd_fill"](ptr(self.owned(self.ring[k])), uniform, self.size, eng.stream))
        else:
            level = self.field._hist.direct_level(t)
            if level is None:
                level = self.field._hist.read({eng.model.time_axis: t})
            eng.dev.upload(level, eng.dtype, out=self.owned(self.ring[k]))
        self.held[k] = t
        self.halo_ok[k] = False
        self.saved.add(t)
        if self.counts is not None:      # NaN count of the level, on the device (FD.py:528)
            self.counts[t] = 0
            nat.check(eng._fn["fd_count_nan"](ptr(self.owned(self.ring[k])), self.size,
                                              self.counts.data_ptr() + 8 * t, eng.stream))

    def level(self, t):
        k = self.slot_index(t)
        if self.constant:
            k = 0
        elif self.held[k] != t:
            raise RuntimeError(
                f"field {self.field.name}: time level {t} is not resident on the device "
                f"(slot holds {self.held[k]})")
        return self.ring[k]

    def claim(self, t):
        """Slot that will receive time level ``t``."""
        k = self.slot_index(t)
        self.engine.wait_for_copies(self, k)
        self.held[k] = t
        self.halo_ok[k] = False
        return self.ring[k]

    def owned(self, tensor):
        """The owned planes of an extended level (all of it without ghosts)."""
        if not self.H:
            return tensor
        return tensor[self.H * self.plane:self.H * self.plane + self.size]

    def refresh_ghosts(self, t):
        """Halo exchange for time level ``t`` (no-op when already current)."""
        k = 0 if self.constant else self.slot_index(t)
        if not self.H:
            return
        self.wants_ghosts = True
        if not self.halo_ok[k]:
            self.engine.slab.exchange(self.ring[k], self.plane, self.owned0, self.H)
            self.halo_ok[k] = True
            self.engine.stats["halo_exchanges"] = self.engine.stats.get("halo_exchanges", 0) + 1
        elif self.halo_event[k] is not None:
            # exchanged ahead of time on the comm stream: order the consumer after it
            if self.halo_event[k] == "seq":        # issued natively (fd_slab_step): its event lives in the library
                nat.check(nat.lib.fd_seq_wait_halo(self.engine.seq, self.engine.stream))
            else:
                self.engine.dev.stream.wait_event(self.halo_event[k])
            self.halo_event[k] = None

    # -- boundary faces ------------------------------------------------------
    def build_faces(self):
        eng, f = self.engine, self.field
        model = eng.model
        space = [a for a in model.axes_names if a != model.time_axis]
        self.faces = {}
        t_from = min(model.time_step + 1, eng.n_t - 1)
        for axis, idx in f._bc_locations:
            dim, side = space.index(axis), (0 if idx == 0 else 1)
            if f._hist.time_dependent({axis: idx}, t_from):
                full = f._hist.read({axis: idx})
                free = [a for a in model.axes_names if a != axis]
                full = self._pad_face(np.moveaxis(full, free.index(model.time_axis), 0), dim, 1)
                per = int(np.prod(full.shape[1:])) if full.ndim > 1 else 1
                self.faces[(dim, side)] = (eng.dev.upload(full, eng.dtype), per)
            else:
                face = self._pad_face(f._hist.read({axis: idx, model.time_axis: t_from}), dim, 0)
                self.faces[(dim, side)] = (eng.dev.upload(face, eng.dtype), 0)

    def _pad_face(self, face, dim, lead):
        """Faces of the other axes span the slab axis: give them the ghost planes
        of the extended layout (NaN = nothing stored there)."""
        if not self.H or dim == 0:
            return face
        pad = [(0, 0)] * face.ndim
        pad[lead] = (self.H, self.H)
        return np.pad(face, pad, constant_values=np.nan)

    def face_ptr(self, dim, side, t):
        if self.faces is None:
            self.build_faces()
        got = self.faces.get((dim, side))
        if got is None:
            return 0
        tensor, per = got
        return tensor.data_ptr() + (t * per * tensor.element_size() if per else 0)

    def face_items(self):
        if self.faces is None:
            self.build_faces()
        return self.faces.items()


class _TempField:
    """Stands where a Field would for the value of an inner expression of a nested
    ``der`` (``d.der(kappa.always * d.der(T.prev, "x"), "x")``): one scratch level,
    no history, no boundary faces."""
    _is_constant = False
    _bc_locations = ()
    _parent = None
    dot = None

    def __init__(self, model, shape, name):
        self.model, self._slice_shape, self.name = model, tuple(shape), name
        self._temp_state = None


class TempState:
    """Device side of a _TempField: what the plan builder asks of a FieldState.
    On a slab it has the layout of the fields (ghost planes on either side of the
    owned planes) so that one descriptor geometry serves both; its ghost planes
    are never filled -- a der along the decomposed axis of a scratch level is
    not fused."""
    constant = integrated = wants_ghosts = False
    counts = peer = pooled = None

    def __init__(self, engine, field):
        self.field, self.engine = field, engine
        shape = tuple(field._slice_shape)
        self.H = engine.H or 0
        self.size = int(np.prod(shape)) if shape else 1
        self.owned0 = shape[0] if shape else 1
        self.plane = self.size // self.owned0 if self.owned0 else 0
        self.ext_shape = ((self.owned0 + 2 * self.H,) + shape[1:]) if self.H else shape
        self.ext_size = self.size + 2 * self.H * self.plane
        self.ring = [engine.dev.empty(self.ext_size, engine.dtype)]
        s = len(shape)
        ext = self.ext_shape
        self.dims3 = (1,) * (3 - s) + ext if s <= 3 else ((int(np.prod(ext[:-2])),) + ext[-2:] if not self.H else None)

    def slot_index(self, t):
        return 0

    def level(self, t):
        return self.ring[0]

    claim = level

    def refresh_ghosts(self, t):
        pass

    def face_items(self):
        return ()

    def face_ptr(self, dim, side, t):
        return 0


class TempLeaf(Leaf):
    """The scratch level of an inner expression, as a source of der terms."""
    __slots__ = ()


class Engine:
    def __init__(self, model):
        self.model = model
        self.dev = Device.get()
        self.dtype = model.dtype
        self.n_t = len(model.axes[model.time_axis])
        self.flags = self.dev.zeros_i64(nat.FD_NUM_FLAGS)
        self.states = {}
        self.slab = model._slab
        self.H = None if self.slab is not None else 0     # ghost depth, fixed at the first executed step
        self.exact_div = os.environ.get("FD_B200_FAST_SCALE", "0") != "1"
        self.fuse = os.environ.get("FD_B200_NO_FUSE", "0") != "1"
        self.split_rim = os.environ.get("FD_B200_NO_SPLIT", "0") != "1"
        self.overlap = os.environ.get("FD_B200_NO_OVERLAP", "0") != "1"
        self.fuse_pairs = os.environ.get("FD_B200_NO_PAIR", "0") != "1"
        self.fuse_expr = os.environ.get("FD_B200_NO_EXPR", "0") != "1"
        self.rims_beside = os.environ.get("FD_B200_NO_SIDE_RIMS", "0") != "1"
        self.fuse_aux = os.environ.get("FD_B200_NO_AUX", "0") != "1"
        self.fuse_nested = os.environ.get("FD_B200_NO_NESTED", "0") != "1"
        # halo planes written straight into the neighbours' memory (one NVLink box, NCCL ranks)
        self.peer_halo = (os.environ.get("FD_B200_NO_P2P", "0") != "1" and self.slab is not None
                          and self.slab.backend == "nccl" and self.dev.comm_stream is not None)
        points = max([int(np.prod(f._slice_shape)) for f in model.fields] or [1])
        default_poll = 16 if points < (1 << 20) else 64
        self.poll_every = int(os.environ.get("FD_B200_POLL_EVERY", default_poll))
        self.last_poll = model.time_step
        self.pending_copies = []          # [(state, slot, t, host tensor, event)]
        self._ghost_readers = {}          # (id(field), level) -> statements of the current step that read its ghost planes
        self._stmt_info = {}              # statement key -> (fields read at the new level, ghost reads relative to t)
        self._events, self._event_next = [], 0     # transient events (recorded and waited for within one call)
        self._halo_done = None            # comm stream: the exchange issued last has completed
        self.seq = None                   # native sequencer of slab steps (fd_seq_create), made on first use
        self.native_seq = os.environ.get("FD_B200_NO_SEQ", "0") != "1"
        self._seq_steps = {}              # statement key -> (fd_slab_step_t, states in link order) or None
        self._slab_fast = {}              # plan key -> (fd_slab_step_t, byref, states, ring slots): the per-step short cut
        self._slab_step_fn = nat.fn("fd_slab_step", self.dtype) if hasattr(nat.lib, "fd_slab_step_f64") else None
        self._ticket = None               # device word of fd_halo_exchange
        self.closed = False
        self._box_cache = {}
        self._plans = {}                  # statement key -> launch descriptor (see _try_fused)
        self.stats = {"fused": 0, "unfused_statements": 0}
        self._fn = {name: nat.fn(name, self.dtype) for name in
                    ("fd_der", "fd_ewise", "fd_merge_store", "fd_fill", "fd_face_store", "fd_count_nan",
                     "fd_fused_step", "fd_fused_pair", "fd_fused_expr", "fd_fused_nested")}

    # ------------------------------------------------------------------ utils
    def state(self, field):
        temp = getattr(field, "_temp_state", None)
        if temp is not None:
            return temp
        st = self.states.get(id(field))
        if st is None:
            if self.H is None:
                raise Exception("on several GPUs expression values can only be read after the first time step")
            st = self.states[id(field)] = FieldState(self, field)
        return st

    @property
    def stream(self):
        return self.dev.stream_ptr

    def _event(self):
        """An event that is recorded and waited for (stream to stream) within the
        calling function: taken round-robin from a small pool instead of being
        created per use.  Re-recording an event does not disturb waits that were
        issued on its earlier record."""
        if len(self._events) < 32:
            self._events.append(self.dev.event())
            return self._events[-1]
        self._event_next = (self._event_next + 1) % len(self._events)
        return self._events[self._event_next]

    def _sequencer(self):
        """The library-side holder of the events (and the ticket word) that native
        multi-launch calls use; made on first use, None on the CPU stand-in."""
        if self.seq is None and self.dev.copy_stream is not None:
            handle = C.c_void_p()
            nat.check(nat.lib.fd_seq_create(C.byref(handle)))
            self.seq = handle
        return self.seq

    def _sole_ghost_reader(self, st, t):
        """True when at most one statement of the current step differentiates
        level ``t`` of this field along the decomposed axis."""
        return self._ghost_readers.get((id(st.field), t), 0) <= 1

    # -------------------------------------------------- rings with peer mappings
    def ring_pool_take(self, st, depth):
        """A ring (with its flag words and neighbour mappings) that an earlier,
        closed model of this process left behind for exactly this geometry."""
        if self.slab is None or not self.peer_halo:
            return None
        from ._dist import RingPool
        return RingPool.take(self._ring_key(st, depth))

    def _ring_key(self, st, depth):
        return (st.ext_size, str(self.dtype), depth, st.plane, st.owned0, st.H, self.slab.world, self.slab.rank,
                self.slab.periodic)

    def close(self):
        """End of the model's life on the device: queued work is finished, rings
        that neighbours have mapped go to the process-wide pool (they must not be
        freed while a peer can still store into them, and the next model of the
        same geometry takes them over together with the mappings)."""
        if self.closed:
            return
        self.closed = True
        try:
            self.dev.synchronize()
        except Exception:
            return
        from ._dist import RingPool
        if self.seq is not None:
            nat.lib.fd_seq_destroy(self.seq)
            self.seq = None
        self._seq_steps = {}
        self._slab_fast = {}
        for st in self.states.values():
            if st.peer:
                st.peer.retire(self.n_t)
                RingPool.give(self._ring_key(st, len(st.ring)),
                              {"ring": st.ring, "flags": st.peer_flags, "links": st.peer})
                st.peer = None
        self.states = {}
        self._plans.clear()
        self._box_cache.clear()
        self.pending_copies = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def wants_level(self, t):
        h = self.model.history
        if h == "last":
            return False          # fetched on demand (Field.data / data_level), nothing streamed
        if t == self.n_t - 1 or h == "all":
            return True
        return t % h == 0

    # --------------------------------------------------------------- execute
    def execute(self, stmts, keys=None):
        """``keys``: the statements' keys when the caller already has them (replay)."""
        if self.closed:
            raise Exception("the model was closed (Model.close): it cannot be advanced any further")
        t = self.model.time_step
        if self.H is None:
            self.H = _slab_reach(stmts)
            if self.H > min(self.slab.counts):
                raise Exception(f"stencils reach {self.H} planes along {self.slab.axis}, more than the "
                                f"{min(self.slab.counts)} planes of the smallest slab")
        if len(self.states) < len(self.model.fields):
            for f in self.model.fields:      # initial levels go up before the first launch
                self.state(f)
        reads_new = set()
        self._ghost_readers = readers = {}
        given = keys
        keys = [None] * len(stmts)
        for i, (_, _, node) in enumerate(stmts):
            if node is not None:
                # what the statement reads depends on its form alone (levels relative to
                # the time cursor): worked out once per form, looked up by its key
                keys[i] = key = given[i] if given is not None else node.key()
                info = self._stmt_info.get(key)
                if info is None:
                    new_reads, ghosts = set(), set()
                    _collect_new_reads(node, t + 1, new_reads)
                    _collect_ghost_reads(node, ghosts)
                    if len(self._stmt_info) >= 256:
                        self._stmt_info.clear()
                    info = self._stmt_info[key] = (frozenset(new_reads), tuple((f, lt - t) for f, lt in ghosts))
                reads_new |= info[0]
                if self.H:
                    for f, rel in info[1]:
                        readers[(f, t + rel)] = readers.get((f, t + rel), 0) + 1
        self.last_keys = keys
        memo = {}
        consumed = set()
        if self.fuse and self.fuse_pairs and self._try_pair(stmts, t, reads_new, keys):
            consumed = set(range(len(stmts)))
            self.stats["fused_pairs"] = self.stats.get("fused_pairs", 0) + 1
        for i, (kind, field, node) in enumerate(stmts):
            if i in consumed:
                continue
            if kind == "assign":
                # (an empty chain: an auxiliary field such as a flux, or a .dot that
                # is not integrated in this step -- the statement is still one launch)
                chain = self._find_chain(stmts, i, field)
                if self.fuse and (chain or self.fuse_aux) and self._try_fused(field, node, chain, t, reads_new, keys[i]):
                    consumed.update(j for j, _ in chain)
                    self.stats["fused"] += 1
                    continue
                self._generic_assign(field, node, t, memo)
            else:
                self._generic_integrate(field, t)
            self.stats["unfused_statements"] += 1
        if self.wants_level(t + 1):
            for f in self.model.fields:
                st = self.state(f)
                if not st.constant:
                    self.snapshot(st, t + 1)

    @staticmethod
    def _find_chain(stmts, i, field):
        chain, parent, start = [], field._parent, i
        while parent is not None and len(chain) < 2:
            found = None
            for j in range(start + 1, len(stmts)):
                if stmts[j][0] == "integrate" and stmts[j][1] is parent:
                    found = j
                    break
            if found is None:
                break
            chain.append((found, parent))
            parent, start = parent._parent, found
        return chain

    # ----------------------------------------------------------- fused path
    def _build_fused_plan(self, dot_field, node, chain, t):
        """Match the statement and fill the launch descriptor for the ring parity of
        step ``t``; returns None when the statement needs the unfused operators."""
        # a der of a computed expression: the inner expression gets a scratch level of
        # its own (one plain-assignment launch), the statement then reads that level
        inner_plans = []
        if _has_nested_der(node):
            if not self.fuse_nested:
                return None
            node, lifted, rewritten = _lift_nested_ders(node, self.model, t)
            if any(der.slab for der in rewritten):
                return None      # the scratch level has no ghost planes: unfused path (which says so)
            for leaf, inner in lifted:
                leaf.field._temp_state = TempState(self, leaf.field)
                plan = self._build_fused_plan(leaf.field, inner, [], t)
                if plan is None:
                    return None
                inner_plans.append(plan)
        dot_state = self.state(dot_field)
        dims3 = dot_state.dims3
        # no chain: the assign_update alone (an auxiliary field such as a flux, or a
        # .dot that is not integrated in this step); its own faces are merged
        aux = not chain
        if dims3 is None or (dot_field._bc_locations and not aux):
            return None
        matched = _match_statement(node)
        program = None
        if matched is None or len(matched[0]) > nat.FD_MAX_TERMS or len(matched[1]) > nat.FD_MAX_POST:
            # not a sum of der terms: a general point-wise expression, evaluated
            # per point by the postfix-program kernel (fd_fused_expr)
            program = _compile_expr(node) if self.fuse_expr else None
            if program is None:
                return None
            terms, posts = [(der, False, None, False) for der in program[0]], []
        else:
            terms, posts = matched
        levels = [self.state(p) for _, p in chain] if chain else [dot_state]
        if any(ls.field._slice_shape != dot_field._slice_shape for ls in levels):
            return None
        if node.shape != tuple(dot_field._slice_shape):
            return None
        nd = len(dot_field._slice_shape)
        pad = 3 - nd
        sdim = max(pad, 0)              # dimension of the kernels' view that holds the decomposed axis
        shape_nd = tuple(dot_field._slice_shape)

        top = nat.Expr() if program is not None else None
        d = top.base if top is not None else nat.Fused()
        sources = []
        for a in range(3):
            d.n[a] = dims3[a]
        d.nterms = len(terms)
        if program is not None:
            _, leaves, code = program
            top.nfields, top.ncode = len(leaves), len(code)
            for j, leaf in enumerate(leaves):
                if leaf.shape != tuple(dot_field._slice_shape):
                    return None
                try:
                    top.field[j] = self.state(leaf.field).level(leaf.t).data_ptr()
                except RuntimeError:
                    return None
                sources.append((self.state(leaf.field), leaf.t - t, False))
            for pc, (op, arg, imm) in enumerate(code):
                top.op[pc], top.arg[pc], top.imm[pc] = op, arg, imm
        for k, (der, subtract, pre, negate) in enumerate(terms):
            leaf = der.a
            if der.shape != tuple(dot_field._slice_shape):
                return None
            src_state = self.state(leaf.field)
            try:
                src = src_state.level(leaf.t)
            except RuntimeError:
                return None
            if max(abs(s) for s in der.shifts) >= der.n_src and der.periodic:
                return None
            tm = d.term[k]
            tm.src = src.data_ptr()
            tm.axis = max(der.axis + pad, 0)
            tm.fold_div = tm.fold_mod = 0
            if nd > 3 and der.axis < nd - 2:
                # one of the merged leading axes (kept extents only: no staggered fixed axis)
                if tuple(leaf.shape) != shape_nd or (not der.slab and der.n_src != shape_nd[der.axis]):
                    return None
                if der.slab and der.n_dst != der.n_src:
                    return None
                tm.fold_mod = src_state.ext_shape[0] if der.slab else shape_nd[der.axis]
                tm.fold_div = int(np.prod(shape_nd[der.axis + 1:nd - 2]))
            if der.slab:
                # the slab is a non-periodic window of the global axis; ghosts hold the neighbours' planes
                tm.periodic = 0
                tm.lo, tm.hi = self._local_range(der, dot_state.owned0)
                tm.n_src = src_state.ext_shape[0]
            else:
                tm.periodic = 1 if der.periodic else 0
                tm.lo, tm.hi, tm.n_src = der.lo, der.hi, der.n_src
            sources.append((src_state, leaf.t - t, bool(der.slab)))
            tm.subtract = 1 if subtract else 0
            tm.has_pre = 0 if pre is None else 1
            tm.pre = 0.0 if pre is None else pre
            tm.negate = 1 if negate else 0
            tm.st = nat.make_stencil(der.shifts, der.weights, der.scale, self.exact_div)
        d.npost = len(posts)
        for k, (op, scalar, leaf) in enumerate(posts):
            d.post[k].op = op
            d.post[k].s = 0.0 if scalar is None else scalar
            if leaf is not None:
                if leaf.shape != tuple(dot_field._slice_shape):
                    return None
                try:
                    d.post[k].coef = self.state(leaf.field).level(leaf.t).data_ptr()
                except RuntimeError:
                    return None
                sources.append((self.state(leaf.field), leaf.t - t, False))
        if any(src_state is dot_state for src_state, _, _ in sources):
            return None          # the target's single slot would be read while it is written
        d.depth = len(levels)
        d.dt = float(self.model.dt)
        dyn_faces = []
        for l, ls in enumerate(levels):
            d.prev[l] = 0 if aux else ls.level(t).data_ptr()      # 0: level 0 is the assigned field itself
            d.out[l] = ls.ring[ls.slot_index(t + 1)].data_ptr()
            for (dim, side), (_, per_time) in ls.face_items():
                if nd > 3 and dim < nd - 2:
                    return None          # a boundary face across merged axes has no place in the 3-D view
                d.face[l][dim + pad][side] = ls.face_ptr(dim, side, t + 1)
                if per_time:
                    dyn_faces.append((l, dim + pad, side, ls, dim))
        d.flags = self.flags.data_ptr()
        for a in range(3):
            d.rim_lo[a], d.rim_hi[a] = 0, max(dims3[a] - 1, 0)
        if self.H:
            unit = dims3[sdim] // dot_state.ext_shape[0]      # merged rows per plane of the decomposed axis (1 up to 3-D)
            d.rim_lo[sdim], d.rim_hi[sdim] = self.H * unit, (self.H + dot_state.owned0) * unit - 1
        for ip in inner_plans:
            sources += ip[3]
        nested = None
        if len(inner_plans) == 1 and not inner_plans[0][9]:
            # the usual case, one inner expression: both launches go through
            # fd_fused_nested, which fills the scratch level in where the
            # descriptors say NULL
            ip = inner_plans[0]
            scratch = ip[1].ring[0]
            inner_d = ip[0]
            inner_d.out[0] = 0
            for a in range(3):
                inner_d.box_lo[a], inner_d.box_hi[a] = 0, ip[6][a]
            if self.H:
                iu = ip[6][ip[5]] // ip[1].ext_shape[0]
                inner_d.box_lo[ip[5]], inner_d.box_hi[ip[5]] = self.H * iu, (self.H + ip[1].owned0) * iu
            hits = 0
            for k in range(d.nterms):
                if d.term[k].src == scratch.data_ptr():
                    d.term[k].src = 0
                    hits += 1
            if top is not None:
                for j in range(top.nfields):
                    if top.field[j] == scratch.data_ptr():
                        top.field[j] = 0
                        hits += 1
            if hits:
                nested = (1 if ip[7] is not None else 0, ip[7] if ip[7] is not None else inner_d, scratch)
        return d, dot_state, levels, sources, dyn_faces, sdim, dims3, top, aux, inner_plans, nested

    def _launch_inner(self, plan):
        """The scratch level of an inner expression (plain assignment, whole level)."""
        d, dot_state, _, _, _, pad, dims3, top, _, inners, _ = plan
        for ip in inners:
            self._launch_inner(ip)
        for a in range(3):
            d.box_lo[a], d.box_hi[a] = 0, dims3[a]
        if self.H:
            unit = dims3[pad] // dot_state.ext_shape[0]
            d.box_lo[pad], d.box_hi[pad] = self.H * unit, (self.H + dot_state.owned0) * unit
        if top is None:
            nat.check(self._fn["fd_fused_step"](C.byref(d), self.stream))
        else:
            nat.check(self._fn["fd_fused_expr"](C.byref(top), self.stream))

    def _try_fused(self, dot_field, node, chain, t, reads_new, node_key=None):
        # the same statement comes back every step: its launch descriptor is
        # built once per ring parity and only the per-step words are patched
        key = (id(dot_field), t & 1, tuple(id(p) for _, p in chain), node_key if node_key is not None else node.key())
        plan = self._plans.get(key)
        if plan is None:
            plan = self._build_fused_plan(dot_field, node, chain, t)
            if plan is None:
                return False
            if len(self._plans) >= 64:      # e.g. a coefficient that changes every step
                self._plans.clear()
                self._box_cache.clear()
                self._seq_steps.clear()
                self._slab_fast.clear()
            self._plans[key] = plan
        d, dot_state, levels, sources, dyn_faces, pad, dims3, top, aux, inner_plans, nested = plan
        for src_state, offset, slab in sources:
            try:
                src_state.level(t + offset)
            except RuntimeError:
                del self._plans[key]
                self._box_cache.clear()
                self._seq_steps.clear()
                self._slab_fast.clear()
                return False
            if slab:
                src_state.refresh_ghosts(t + offset)
        if aux:
            d.dot_out = 0
            d.out[0] = dot_state.claim(t + 1).data_ptr()
        else:
            want_dot = id(dot_field) in reads_new or self.wants_level(t + 1)
            d.dot_out = dot_state.claim(t + 1).data_ptr() if want_dot else 0
            if not want_dot:
                dot_state.held[0] = -1
            for l, ls in enumerate(levels):
                ls.level(t)
                ls.claim(t + 1)
                d.nan_count[l] = ls.counts.data_ptr() + 8 * (t + 1)
        for l, dim3, side, ls, dim in dyn_faces:
            d.face[l][dim3][side] = ls.face_ptr(dim, side, t + 1)
        fast = self._slab_fast.get(key)
        if fast is not None and not all(self._sole_ghost_reader(st, t) for st in fast[2]):
            fast = None          # another statement of this step reads the same ghost planes: the general path decides
        if fast is not None:
            # a slab statement that went through fd_slab_step before: its step
            # descriptor is complete, only the level number of the exchange changes
            step, ref, states, slots = fast
            for j, st in enumerate(states):
                step.link[j].value = st.peer.base + t + 2
            step.compute_stream = self.stream
            nat.check(self._slab_step_fn(self.seq, ref))
            for st, k in zip(states, slots):
                st.halo_ok[k], st.halo_event[k] = True, "seq"
            n, stats = len(states), self.stats
            stats["halo_exchanges"] = stats.get("halo_exchanges", 0) + n
            stats["peer_halo_exchanges"] = stats.get("peer_halo_exchanges", 0) + n
            stats["native_slab_steps"] = stats.get("native_slab_steps", 0) + 1
            return True
        region = [[0, dims3[a]] for a in range(3)]
        unit = 1
        if self.H:
            # only the owned planes are computed; the domain faces of the slab
            # axis sit on the first / last owned plane of the end ranks
            unit = dims3[pad] // dot_state.ext_shape[0]
            region[pad] = [self.H * unit, (self.H + dot_state.owned0) * unit]
        launch = self._fn["fd_fused_step" if top is None else "fd_fused_expr"]
        ref = C.byref(d if top is None else top)
        if top is not None:
            self.stats["fused_expr"] = self.stats.get("fused_expr", 0) + 1
        if inner_plans:
            self.stats["fused_nested"] = self.stats.get("fused_nested", 0) + 1
            for a in range(3):
                d.box_lo[a], d.box_hi[a] = region[a]
            if nested is not None:
                inner_is_expr, inner_desc, scratch = nested
                nat.check(self._fn["fd_fused_nested"](inner_is_expr, C.addressof(inner_desc), 0 if top is None else 1,
                                                      C.addressof(d if top is None else top), scratch.data_ptr(),
                                                      scratch.numel(), self.stream))
            else:
                for ip in inner_plans:
                    self._launch_inner(ip)
                nat.check(launch(ref, self.stream))
            return True

        def run(piece, on=None):
            base = on if on is not None else self.dev.stream        # None on the CPU stand-in
            base_ptr = base.cuda_stream if on is not None else self.stream
            bkey = (id(d), tuple(map(tuple, piece)))
            boxes = self._box_cache.get(bkey)
            if boxes is None:
                # (the expression kernel has a single instantiation: nothing to gain from an interior box)
                boxes = self._box_cache[bkey] = self._split_boxes(d, piece) if top is None else [piece]
            rim_stream = self.dev.rim_stream if len(boxes) > 1 else None
            if rim_stream is not None and top is None and self.native_seq and self._sequencer() is not None:
                # interior box + rim slabs, their two events and two stream waits in one native call
                packed = self._box_cache.get(("packed",) + bkey)
                if packed is None:
                    packed = self._box_cache[("packed",) + bkey] = (C.c_int32 * (6 * len(boxes)))(
                        *[v for box in boxes for a in range(3) for v in box[a]])
                nat.check(nat.fn("fd_fused_step_boxes", self.dtype)(self.seq, C.addressof(d), len(boxes), packed, base_ptr,
                                                                    rim_stream.cuda_stream))
                return
            if rim_stream is not None:
                # the rim slabs read the same level and write disjoint points:
                # they run beside the interior box instead of after it
                start = self._event()
                start.record(base)
                rim_stream.wait_event(start)
            for i, box in enumerate(boxes):
                for a in range(3):
                    d.box_lo[a], d.box_hi[a] = box[a]
                stream = rim_stream.cuda_stream if (i and rim_stream is not None) else base_ptr
                nat.check(launch(ref, stream))
            if rim_stream is not None:
                done = self._event()
                done.record(rim_stream)
                base.wait_event(done)

        # (decided from facts every rank shares: slabs differ in size, collectives must not)
        eager = [ls for ls in levels if ls.H and ls.wants_ghosts and min(self.slab.counts) >= 3 * ls.H]
        if eager and self.overlap:
            # peer-memory pushes need the statement itself to read the level-t ghosts
            # of the field it advances (see _exchange_ahead)
            own = {id(st) for st, offset, slab in sources if slab and offset == 0}
            self._run_overlapped(run, region, pad, eager, t,
                                 {id(ls) for ls in eager if id(ls) in own and not aux and self._sole_ghost_reader(ls, t)},
                                 seq_desc=(0, d, d, key) if top is None else (2, top, d, key), depth=self.H * unit)
        else:
            run(region)
        return True

    def _native_slab_step(self, seq_desc, region, pad, eager, t, direct_ok, depth):
        """The whole of _run_overlapped in one native call (fd_slab_step) when every
        exchanged field travels through peer memory and no piece needs an
        interior/rim split; False if the statement has to take the Python path."""
        kind, d, base_desc, cache_key = seq_desc
        dev = self.dev
        if not (self.native_seq and self.peer_halo and self.rims_beside and dev.comm_stream is not None):
            return False
        for st in eager:                    # collective set-up, first use only
            if st.peer is None and id(st) in direct_ok and len(st.ring) > 1:
                st.peer = self._peer_links(st)
        if not all(st.peer and id(st) in direct_ok for st in eager) or len(eager) > nat.FD_MAX_LINKS:
            return False
        cached = self._seq_steps.get(cache_key, 0)
        if cached == 0:
            cached = None
            lo, hi = region[pad]
            pieces = [[lo, lo + depth], [hi - depth, hi], [lo + depth, hi - depth]]
            whole = lambda cut: [list(r) if a != pad else cut for a, r in enumerate(region)]
            single = kind != 0 or all(len(self._split_boxes(base_desc, whole(c))) == 1 for c in pieces)
            if single:
                step = nat.SlabStep()
                step.kind, step.dim, step.lo, step.hi, step.depth = kind, pad, lo, hi, depth
                step.nlinks = len(eager)
                step.timeout_ns = 20_000_000_000
                step.flags = self.flags.data_ptr()
                cached = (step, list(eager))
            self._seq_steps[cache_key] = cached
        if cached is None:
            return False
        step, states = cached
        if [id(s) for s in states] != [id(s) for s in eager]:
            return False
        self._sequencer()
        for a in range(3):
            base_desc.box_lo[a], base_desc.box_hi[a] = region[a]
        for j, st in enumerate(eager):
            st.peer.fill(step.link[j], st.slot_index(t + 1), t + 2)
        step.desc = C.addressof(d)
        step.compute_stream, step.comm_stream = self.stream, dev.comm_stream.cuda_stream
        nat.check(nat.fn("fd_slab_step", self.dtype)(self.seq, C.byref(step)))
        if kind != 1 and top_level_plan_key(cache_key):
            # (fused / expression statements have a cached plan whose descriptor lives on:
            # from the next step of this ring parity on _try_fused calls the library directly)
            self._slab_fast[cache_key] = (step, C.byref(step), list(eager), [st.slot_index(t + 1) for st in eager])
        for st in eager:
            k = st.slot_index(t + 1)
            st.halo_ok[k], st.halo_event[k] = True, "seq"
        self.stats["halo_exchanges"] = self.stats.get("halo_exchanges", 0) + len(eager)
        self.stats["peer_halo_exchanges"] = self.stats.get("peer_halo_exchanges", 0) + len(eager)
        self.stats["native_slab_steps"] = self.stats.get("native_slab_steps", 0) + 1
        return True

    def _run_overlapped(self, run, region, pad, eager, t, direct_ok=(), seq_desc=None, depth=None):
        """One statement (group) of a slab: the boundary planes and their exchange
        on the high-priority comm stream, the interior beside them on the compute
        stream, so that neither the thin boundary launches nor the trip over
        NVLink add to the step."""
        lo, hi = region[pad]
        depth = depth or self.H              # ghost depth in rows of the kernels' view (planes up to 3-D)
        piece = lambda cut: [list(r) if a != pad else cut for a, r in enumerate(region)]
        cuts = ([lo, lo + depth], [hi - depth, hi])
        inner = [lo + depth, hi - depth]
        dev = self.dev
        if seq_desc is not None and self._native_slab_step(seq_desc, region, pad, eager, t, direct_ok, depth):
            return
        if dev.comm_stream is None or not self.rims_beside:
            # (CPU stand-in, or FD_B200_NO_SIDE_RIMS=1) boundary planes first, their
            # exchange on the comm stream, then the interior while the planes travel
            for cut in cuts:
                run(piece(cut))
            self._exchange_ahead(eager, t + 1, direct_ok=direct_ok)
            run(piece(inner))
            return
        start = self._event()
        start.record(dev.stream)
        # the interior is queued first: when the GPU is idle (first step after a
        # synchronisation) it starts at once instead of after the host has issued
        # the boundary launches and the exchange; those go to the high-priority
        # stream and get the next free SMs.  In steady state the host runs ahead
        # and the order of issue does not matter.
        run(piece(inner))
        dev.comm_stream.wait_event(start)
        for cut in cuts:
            run(piece(cut), dev.comm_stream)
        rims_done = self._event()
        rims_done.record(dev.comm_stream)
        self._exchange_ahead(eager, t + 1, on_comm=True, direct_ok=direct_ok)
        dev.stream.wait_event(rims_done)      # later work on the compute stream sees the whole level

    def _exchange_ahead(self, states, t, on_comm=False, direct_ok=()):
        """Halo exchange of freshly written boundary planes on the comm stream
        (on_comm: the planes were computed on that stream, nothing to wait for).

        direct_ok: states that may be written straight into the neighbours' ghost
        planes.  A push is not a rendezvous: nothing stops a neighbour from
        delivering level t+2 into the slot whose level-t ghosts this rank is
        still reading -- unless the statement that produces a level also reads
        the previous level's ghosts of the same field.  Then a neighbour can only
        produce (and push) level t+2 after it has received this rank's level t+1
        planes, which are pushed after the boundary kernels that read the level-t
        ghosts.  Heat/wave statements and the staggered pair have that form; any
        other field is exchanged with isend/irecv, which is a rendezvous."""
        dev = self.dev
        if dev.comm_stream is None:          # CPU stand-in: same order of operations, synchronously
            for st in states:
                k = st.slot_index(t)
                self.slab.exchange(st.ring[k], st.plane, st.owned0, st.H)
                st.halo_ok[k], st.halo_event[k] = True, None
                self.stats["halo_exchanges"] = self.stats.get("halo_exchanges", 0) + 1
            return
        import torch
        if self.peer_halo:
            for st in states:               # collective set-up, first use only
                if st.peer is None and id(st) in direct_ok and len(st.ring) > 1:
                    st.peer = self._peer_links(st)
        if not on_comm:
            ready = self._event()
            ready.record(dev.stream)
        with torch.cuda.stream(dev.comm_stream):
            if not on_comm:
                dev.comm_stream.wait_event(ready)
            cs = dev.comm_stream.cuda_stream
            direct = [st for st in states if st.peer and id(st) in direct_ok][:nat.FD_MAX_LINKS]
            # peer memory: copy, publication and wait of all fields in one launch
            if direct:
                if self._ticket is None:
                    self._ticket = dev.zeros_i64(1)
                links = (nat.HaloLink * len(direct))()
                for j, st in enumerate(direct):
                    st.peer.fill(links[j], st.slot_index(t), t + 1)
                nat.check(nat.lib.fd_halo_exchange(links, len(direct), self._ticket.data_ptr(), 20_000_000_000,
                                                   self.flags.data_ptr(), cs))
            for st in states:
                if st not in direct:
                    k = st.slot_index(t)
                    self.slab.exchange(st.ring[k], st.plane, st.owned0, st.H)
            if direct:
                self.stats["peer_halo_exchanges"] = self.stats.get("peer_halo_exchanges", 0) + len(direct)
            # one event, always recorded on the comm stream: a consumer that waits
            # for it after a later exchange was recorded waits for that later point
            # of the same stream, which lies behind the one it needs
            if self._halo_done is None:
                self._halo_done = dev.event()
            done = self._halo_done
            done.record(dev.comm_stream)
        for st in states:
            k = st.slot_index(t)
            st.halo_ok[k], st.halo_event[k] = True, done
            self.stats["halo_exchanges"] = self.stats.get("halo_exchanges", 0) + 1

    def _peer_links(self, st):
        """Map the neighbours' copies of this field's ring (collective); False if
        any rank could not, in which case the field keeps isend/irecv.  A ring
        taken over from a closed model keeps its mappings when every rank has one."""
        from ._dist import PeerLinks
        pooled = st.pooled
        st.pooled = None
        reuse = self.slab.all_reduce(self.dev.ones_i64(1) if pooled is not None else self.dev.zeros_i64(1), "min")
        if int(reuse.item()):
            st.peer_flags = pooled["flags"]
            self.stats["peer_links_reused"] = self.stats.get("peer_links_reused", 0) + 1
            return pooled["links"]
        # fresh mappings: flag words start from zero on every rank (the collectives
        # inside PeerLinks separate this reset from the first push of any rank)
        st.peer_flags = self.dev.zeros_i64(2 * len(st.ring))
        self.dev.synchronize()
        links = PeerLinks(self.slab, nat, st.ring, st.peer_flags, st.plane * self.dtype.itemsize, st.owned0, st.H)
        return links if links.ok else False

    def _split_boxes(self, d, region):
        """Interior box first, then the rim slabs around it.

        In the interior no output falls outside a fixed axis' [lo, hi) and no
        boundary face applies, so the kernel runs its predicate-free
        instantiation there; the thin rim slabs take the bounded one.  Every
        point is still computed exactly once, in one pass."""
        if not self.split_rim or any(d.term[k].fold_mod for k in range(d.nterms)):
            return [region]
        inner = [list(r) for r in region]
        for k in range(d.nterms):
            tm = d.term[k]
            if not tm.periodic:
                inner[tm.axis][0] = max(inner[tm.axis][0], tm.lo)
                inner[tm.axis][1] = min(inner[tm.axis][1], tm.hi)
        for l in range(d.depth):
            for a in range(3):
                if d.face[l][a][0]:
                    inner[a][0] = max(inner[a][0], d.rim_lo[a] + 1)
                if d.face[l][a][1]:
                    inner[a][1] = min(inner[a][1], d.rim_hi[a])
        if inner == region:
            return [region]
        # the contiguous axis is processed in whole vector groups
        align = 64 * (16 // self.dtype.itemsize)
        # 3-point statements: the kernels can handle a fixed contiguous axis
        # themselves; worth it where a z rim of whole vector groups would be a
        # large share of the row (2 x 128 of 512 points), not on long rows (8192)
        if (all(d.term[k].st.npts == 3 for k in range(d.nterms)) and d.depth == 1
                and 64 * align > region[2][1] - region[2][0]):
            inner[2] = list(region[2])
        if inner[2] != region[2]:
            lo = -(-inner[2][0] // align) * align
            hi = inner[2][1] // align * align
            inner[2] = [lo, hi]
        vol = lambda b: max(0, b[0][1] - b[0][0]) * max(0, b[1][1] - b[1][0]) * max(0, b[2][1] - b[2][0])
        if vol(inner) * 2 < vol(region):
            return [region]
        boxes = [inner]
        cur = [list(r) for r in region]
        for a in range(3):                       # peel the rim off axis by axis
            if inner[a][0] > cur[a][0]:
                boxes.append([list(c) if i != a else [cur[a][0], inner[a][0]] for i, c in enumerate(cur)])
            if inner[a][1] < cur[a][1]:
                boxes.append([list(c) if i != a else [inner[a][1], cur[a][1]] for i, c in enumerate(cur)])
            cur[a] = list(inner[a])
        return boxes

    def _local_range(self, der, owned_out):
        """[lo, hi) of a der along the slab axis in extended local plane indices:
        the global range (FD.py:1073-1105) cut to the planes this rank owns."""
        reach = max(abs(s) for s in der.shifts)
        if reach > self.H:
            raise Exception(f"a stencil reaches {reach} planes along the decomposed axis {self.slab.axis}, but the "
                            f"slabs were laid out with {self.H} ghost planes (the reach seen in the first time step)")
        lo_g, hi_g = (0, der.n_dst) if der.periodic else (der.lo, der.hi)
        first = self.slab.lo
        lo = max(lo_g, first) - first + self.H
        hi = min(hi_g, first + owned_out) - first + self.H
        return lo, max(hi, lo)


    # ------------------------------------------------- fused staggered pair
    def _try_pair(self, stmts, t, reads_new, keys=None):
        """The whole step is  v_a += dt*c_a*up.der(p.prev, a)  for some axes a, then
        p += dt*c*(sum_a down.der(v_a.new, a))  (cell/edge leapfrog):
        one launch of fd_fused_pair instead of len(axes)+1 fused statements.
        The match is made once per form of the step and ring parity; afterwards
        only the per-step words of the descriptor are patched."""
        n = len(stmts)
        if n < 4 or n % 2 or n > 8:
            return False
        pkey = ("pair", t & 1, tuple(id(f) for _, f, _ in stmts),
                tuple(keys) if keys is not None else tuple(None if nd is None else nd.key() for _, _, nd in stmts))
        plan = self._plans.get(pkey, 0)
        if plan == 0:
            plan = self._build_pair_plan(stmts, t)
            if len(self._plans) >= 64:
                self._plans.clear()
                self._box_cache.clear()
                self._seq_steps.clear()
                self._slab_fast.clear()
            self._plans[pkey] = plan          # (None: this form of step is not a staggered pair)
        if plan is None:
            return False
        d, p_field, p_state, pd_state, axes, region, pad = plan
        try:
            if self.H:
                p_state.refresh_ghosts(t)
            d.p_prev = p_state.level(t).data_ptr()
            wants = self.wants_level(t + 1)
            outs = []
            for ax, vfield, vs, vds, along_slab in axes:
                if along_slab:
                    vs.refresh_ghosts(t)
                ax.v_prev = vs.level(t).data_ptr()
        except RuntimeError:                  # a level is not resident: the general paths report it
            del self._plans[pkey]
            return False
        for ax, vfield, vs, vds, along_slab in axes:
            ax.v_out = vs.claim(t + 1).data_ptr()
            ax.v_nan_count = vs.counts.data_ptr() + 8 * (t + 1)
            want = wants or id(vfield.dot) in reads_new
            ax.vdot_out = vds.claim(t + 1).data_ptr() if want else 0
            if not want:
                vds.held[0] = -1
            outs.append(vs)
        d.p_out = p_state.claim(t + 1).data_ptr()
        d.p_nan_count = p_state.counts.data_ptr() + 8 * (t + 1)
        want = wants or id(p_field.dot) in reads_new
        d.pdot_out = pd_state.claim(t + 1).data_ptr() if want else 0
        if not want:
            pd_state.held[0] = -1
        launch = self._fn["fd_fused_pair"]

        def run(piece, on=None):
            for a in range(3):
                d.box_lo[a], d.box_hi[a] = piece[a]
            nat.check(launch(C.byref(d), on.cuda_stream if on is not None else self.stream))

        eager = [st for st in [p_state] + outs if st.H and st.wants_ghosts and min(self.slab.counts) >= 3 * st.H]
        if eager and self.overlap:
            # (the pair kernel reads the level-t ghosts of p and of the slab-axis velocity,
            # and the statements it replaces are the only ones of the step)
            self._run_overlapped(run, region, pad, eager, t, {id(st) for st in eager},
                                 seq_desc=(1, d, d, ("pair", t & 1, len(axes))))
        else:
            run(region)
        return True

    def _build_pair_plan(self, stmts, t):
        """Match the statements of a step against the staggered pair and fill
        everything of fd_pair_t that does not change from step to step; None if
        the step has another form."""
        n = len(stmts)
        kinds = [k for k, _, _ in stmts]
        if kinds != ["assign", "integrate"] * (n // 2):
            return None
        for i in range(0, n, 2):                       # assign(F.dot) directly followed by integrate(F)
            if stmts[i][1]._parent is not stmts[i + 1][1] or stmts[i + 1][1]._parent is not None:
                return None
        p_field = stmts[n - 1][1]
        p_state = self.state(p_field)
        dims3 = p_state.dims3
        vn = 16 // self.dtype.itemsize
        if dims3 is None or dims3[2] % vn or dims3[2] < 3 * vn or len(p_field._slice_shape) > 3:
            return None
        shape = tuple(p_field._slice_shape)
        pad = 3 - len(shape)
        down = _match_statement(stmts[n - 2][2])
        if down is None:
            return None
        terms, posts = down
        if len(terms) != n // 2 - 1 or posts not in ([], [(nat.POST_MUL, posts[0][1] if posts else 0.0, None)]):
            return None
        ups = {}
        for i in range(0, n - 2, 2):                   # the velocity statements
            m = _match_statement(stmts[i][2])
            if m is None:
                return None
            (uterms, uposts), vfield = m, stmts[i + 1][1]
            if len(uterms) != 1 or uposts not in ([], [(nat.POST_MUL, uposts[0][1] if uposts else 0.0, None)]):
                return None
            der, _, pre, negate = uterms[0]
            if (negate or (pre is not None and uposts) or der.shifts != (-1, 0) or not isinstance(der.a, Leaf)
                    or der.a.field is not p_field or der.a.t != t or tuple(vfield._slice_shape) != shape):
                return None
            # `c * der` is matched as a pre-scaled term; as a post-multiplication it rounds the same
            c_up = pre if pre is not None else (uposts[0][1] if uposts else 1.0)
            ups[id(vfield)] = (vfield, der, c_up)
        fields = [p_field, p_field.dot] + [f for f, _, _ in ups.values()] + [f.dot for f, _, _ in ups.values()]
        if any(f._bc_locations for f in fields) or len(ups) != len(terms):
            return None
        d = nat.Pair()
        used = set()
        axes = []
        for k, (der, subtract, pre, negate) in enumerate(terms):
            leaf = der.a
            got = ups.get(id(leaf.field)) if isinstance(leaf, Leaf) else None
            if (got is None or leaf.t != t + 1 or pre is not None or negate or der.shifts != (0, 1)
                    or got[1].axis != der.axis or der.axis in used):
                return None
            used.add(der.axis)
            vfield, uder, c_up = got
            slab_axis = bool(self.H) and der.axis == 0
            # (the ders' own flags are global facts, the same on every rank: a fixed
            # axis needs the range predicates of the general kernels, also along the slabs)
            if not der.periodic or not uder.periodic:
                return None
            if slab_axis and self.H != 1:
                return None
            ax = d.ax[k]
            ax.axis, ax.periodic, ax.subtract = der.axis + pad, 0 if slab_axis else 1, 1 if subtract else 0
            ax.c_up = c_up
            ax.up = nat.make_stencil(uder.shifts, uder.weights, uder.scale, self.exact_div)
            ax.down = nat.make_stencil(der.shifts, der.weights, der.scale, self.exact_div)
            axes.append((ax, vfield, self.state(vfield), self.state(vfield.dot), slab_axis))
        # every axis of the grid that is not differentiated must not matter; every
        # differentiated one is periodic or the slab axis: nothing can be NaN-masked
        if self.H and 0 not in used:
            return None
        for a in range(3):
            d.n[a] = dims3[a]
        d.naxes = len(terms)
        d.c_down = posts[0][1] if posts else 1.0
        d.dt = float(self.model.dt)
        d.flags = self.flags.data_ptr()
        region = [[0, dims3[a]] for a in range(3)]
        if self.H:
            region[pad] = [self.H, self.H + p_state.owned0]
        return d, p_field, p_state, self.state(p_field.dot), axes, region, pad

    # --------------------------------------------------------- unfused path
    def _eval(self, node, memo):
        """Evaluate an expression node with one launch per reference operation."""
        got = memo.get(id(node))
        if got is not None:
            return got
        n = self._ext_elems(node.shape)
        if isinstance(node, Leaf):
            out = self.state(node.field).level(node.t)
        elif isinstance(node, Const):
            if self.H:
                raise Exception("user-built FieldInstant arrays are not supported on several GPUs")
            out = self.dev.upload(node.array, self.dtype)
        elif isinstance(node, Unary):
            a = self._eval(node.a, memo)
            out = self.dev.empty(n, self.dtype)
            nat.check(self._fn["fd_ewise"](nat.OP_NEG, ptr(out), ptr(a), 0, 0.0, n, self.stream))
        elif isinstance(node, Binary):
            out = self._eval_binary(node, n, memo)
        elif isinstance(node, Der):
            a = self._eval(node.a, memo)
            shape = self._ext_shape(node.a.shape)
            outer = int(np.prod(shape[:node.axis])) if node.axis else 1
            inner = int(np.prod(shape[node.axis + 1:])) if node.axis + 1 < len(shape) else 1
            st = nat.make_stencil(node.shifts, node.weights, node.scale, self.exact_div)
            out = self.dev.empty(n, self.dtype)
            n_src, n_dst, periodic, lo, hi = node.n_src, node.n_dst, node.periodic, node.lo, node.hi
            if node.slab:
                if not isinstance(node.a, Leaf):
                    raise NotImplementedError(
                        "on several GPUs a derivative along the decomposed axis must be taken of a stored "
                        "field level (prev/new/always), not of an intermediate expression")
                self.state(node.a.field).refresh_ghosts(node.a.t)
                n_src, n_dst, periodic = shape[0], node.shape[0] + 2 * self.H, False
                lo, hi = self._local_range(node, node.shape[0])
            nat.check(self._fn["fd_der"](ptr(out), ptr(a), outer, n_src, n_dst, inner, C.byref(st),
                                          1 if periodic else 0, lo, hi, ptr(self.flags), self.stream))
        else:
            raise TypeError(f"unknown expression node {type(node).__name__}")
        memo[id(node)] = out
        return out

    def _ext_shape(self, shape):
        shape = tuple(shape)
        return ((shape[0] + 2 * self.H,) + shape[1:]) if (self.H and shape) else shape

    def _ext_elems(self, shape):
        shape = self._ext_shape(shape)
        return int(np.prod(shape)) if shape else 1

    def _eval_binary(self, node, n, memo):
        out = self.dev.empty(n, self.dtype)
        ew = self._fn["fd_ewise"]
        if isinstance(node.b, Node) and isinstance(node.a, Node):
            if node.a.shape != node.b.shape:
                raise ValueError(f"operands could not be broadcast together with shapes {node.a.shape} {node.b.shape}")
            a, b = self._eval(node.a, memo), self._eval(node.b, memo)
            nat.check(ew(_FF_OPS[node.op], ptr(out), ptr(a), ptr(b), 0.0, n, self.stream))
            return out
        a, s = self._eval(node.a, memo), float(node.b)
        if node.op == "pow":
            # numpy's scalar-exponent fast paths (square, sqrt, reciprocal, copy)
            op = {2.0: nat.OP_SQUARE, 0.5: nat.OP_SQRT, -1.0: nat.OP_RECIP, 1.0: nat.OP_COPY}.get(s, nat.OP_POW_FS)
        else:
            op = _FS_OPS[node.op]
        nat.check(ew(op, ptr(out), ptr(a), 0, s, n, self.stream))
        return out

    def _prefilled(self, st, t):
        """Claim the slot for level t and give it the reference's pre-populated
        content: NaN plus the boundary faces of that time (FD.py:310, 361-385)."""
        slot = st.claim(t)
        nat.check(self._fn["fd_fill"](ptr(slot), float("nan"), st.ext_size, self.stream))
        shape = st.ext_shape
        for (dim, side), _ in st.face_items():
            outer = int(np.prod(shape[:dim])) if dim else 1
            inner = int(np.prod(shape[dim + 1:])) if dim + 1 < len(shape) else 1
            index = 0 if side == 0 else shape[dim] - 1
            if dim == 0 and st.H:
                index = st.H if side == 0 else st.H + st.owned0 - 1
            nat.check(self._fn["fd_face_store"](ptr(slot), st.face_ptr(dim, side, t), outer, shape[dim], inner,
                                                 index, self.stream))
        return slot

    def _generic_assign(self, field, node, t, memo):
        if node.shape != tuple(field._slice_shape):
            raise ValueError(f"could not broadcast input array from shape {node.shape} into shape {tuple(field._slice_shape)}")
        st = self.state(field)
        fresh = self._eval(node, memo)
        slot = self._prefilled(st, t + 1)
        nat.check(self._fn["fd_merge_store"](ptr(st.owned(slot)), ptr(st.owned(fresh)), st.size, 0,
                                              ptr(self.flags), self.stream))

    def _generic_integrate(self, field, t):
        st, dot = self.state(field), self.state(field.dot)
        fresh = self.dev.empty(st.ext_size, self.dtype)
        nat.check(self._fn["fd_ewise"](nat.OP_EULER, ptr(fresh), ptr(st.level(t)), ptr(dot.level(t + 1)),
                                        float(self.model.dt), st.ext_size, self.stream))
        slot = self._prefilled(st, t + 1)
        nat.check(self._fn["fd_merge_store"](ptr(st.owned(slot)), ptr(st.owned(fresh)), st.size,
                                              st.counts.data_ptr() + 8 * (t + 1), ptr(self.flags), self.stream))

    # ------------------------------------------------------ host interaction
    def evaluate_to_host(self, node):
        out = self._eval(node, {})
        if self.H:
            plane = int(np.prod(node.shape[1:])) if len(node.shape) > 1 else 1
            out = out[self.H * plane:(self.H + node.shape[0]) * plane]
        arr = self.dev.download(out, node.shape)
        self.poll(final=True)
        return np.array(arr, dtype=np.float64)

    def snapshot(self, st, t):
        """Device level -> pinned host memory, asynchronously on the copy stream.

        The compute stream only records an event; the copy stream waits for it,
        copies the ring slot into pinned memory and records a second event that
        the compute stream waits on before it overwrites that slot again
        (``wait_for_copies``)."""
        k = st.slot_index(t)
        if t in st.saved or st.held[k] != t:
            return
        st.saved.add(t)
        host = self.dev.pinned(st.size, self.dtype)
        if self.dev.copy_stream is None:              # CPU stand-in used by the host-logic tests
            host.copy_(st.owned(st.ring[k]))
            self.pending_copies.append((st, k, t, host, None))
            return
        tc = self.dev
        import torch
        ready = self._event()
        ready.record(tc.stream)
        with torch.cuda.stream(tc.copy_stream):
            tc.copy_stream.wait_event(ready)
            host.copy_(st.owned(st.ring[k]), non_blocking=True)
            done = tc.event()
            done.record(tc.copy_stream)
        self.pending_copies.append((st, k, t, host, done))
        if len(self.pending_copies) > 64:
            self.drain_copies(keep=32)

    def fetch_into(self, field, t, out):
        """Level ``t`` of ``field`` from its ring slot into the caller's host array."""
        st = self.states.get(id(field))
        if st is None or st.constant or st.held[st.slot_index(t)] != t:
            return False
        if out.shape != tuple(field._slice_shape) or out.dtype != self.dtype or not out.flags.c_contiguous:
            return False
        import torch
        torch.from_numpy(out.reshape(-1)).copy_(st.owned(st.ring[st.slot_index(t)]))
        self.poll(final=True)
        return True

    def wait_for_copies(self, st, slot):
        """The compute stream must not overwrite a ring slot that is still being
        copied out."""
        for (pst, k, _, _, done) in self.pending_copies:
            if pst is st and k == slot and done is not None:
                self.dev.stream.wait_event(done)

    def drain_copies(self, field=None, keep=0):
        """Hand finished (or, with keep=0, all) snapshot copies to the host history."""
        rest = []
        n = len(self.pending_copies)
        for i, item in enumerate(self.pending_copies):
            st, k, t, host, done = item
            if (field is not None and st.field is not field) or i >= n - keep:
                rest.append(item)
                continue
            if done is not None:
                done.synchronize()
            arr = host.numpy().reshape(st.field._slice_shape)
            if arr.dtype != np.float64:
                arr = arr.astype(np.float64)
            st.field._hist.write_level(t, arr)
        self.pending_copies = rest

    def sync_history(self, field):
        """Bring ``Field.data`` up to date: outstanding copies plus the level the
        model currently stands on."""
        st = self.states.get(id(field))
        if st is None or st.constant:
            return
        self.snapshot(st, self.model.time_step)
        self.drain_copies(field)
        self.poll(final=True)

    def host_changed(self, field):
        self._plans.clear()
        self._box_cache.clear()
        self._seq_steps.clear()
        self._slab_fast.clear()
        st = self.states.get(id(field))
        if st is not None:
            st.faces = None
            if st.constant:
                st.load_level(0)

    # ---------------------------------------------------------------- polling
    def step_closed(self):
        if self.model.time_step - self.last_poll >= self.poll_every:
            self.poll()

    def poll(self, final=False):
        """Read the sticky device flags and NaN counters; raise/print what the
        reference would have raised/printed during the steps since the last poll."""
        now = self.model.time_step
        if now == self.last_poll and not final:
            return
        self.last_poll = now
        if self.slab is not None:            # every rank polls at the same steps
            self.slab.all_reduce(self.flags, "max")
        flags = self.flags.cpu().numpy()    # synchronises the compute stream
        if flags[nat.FD_FLAG_HALO_TIMEOUT]:
            raise Exception("halo exchange: a neighbouring rank did not deliver its boundary planes in time")
        if flags[nat.FD_FLAG_INF] or flags[nat.FD_FLAG_OVERRIDE]:
            self.flags.zero_()
            if flags[nat.FD_FLAG_INF]:
                raise Exception(INSTABILITY_MSG)
            raise Exception(OVERRIDE_MSG)
        if self.model.insufficient_bc:
            return
        for st in self.states.values():
            if st.counts is None or st.counts_checked >= now:
                continue
            window = st.counts[st.counts_checked:now + 1]
            if self.slab is not None:
                window = self.slab.all_reduce(window.clone(), "sum")
            counts = window.cpu().numpy()
            first = st.counts_checked
            st.counts_checked = now
            for k in range(1, len(counts)):
                # FD.py:535: only from the second step on, and only once per model
                if first + k > 1 and counts[k] > counts[k - 1]:
                    print(LOSING_MSG)
                    self.model.insufficient_bc = True
                    return


_X_BIN = {"add": nat.X_ADD, "sub": nat.X_SUB, "mul": nat.X_MUL, "truediv": nat.X_DIV, "pow": nat.X_POW}
_X_IMM = {"add": nat.X_ADD_IMM, "sub": nat.X_SUB_IMM, "mul": nat.X_MUL_IMM, "truediv": nat.X_DIV_IMM}
_X_POW = {2.0: nat.X_SQUARE, 0.5: nat.X_SQRT, -1.0: nat.X_RECIP}   # numpy's scalar-exponent fast paths


def _compile_expr(node):
    """Expression tree -> postfix program for fd_fused_expr: (der nodes, field
    leaves, [(op, arg, imm)]), or None if it does not fit the kernel's limits
    (der of an intermediate expression, too many operands, too deep)."""
    ders, leaves, code = [], [], []
    depth = [0, 0]

    def push(n=1):
        depth[0] += n
        depth[1] = max(depth[1], depth[0])

    def emit(n):
        if isinstance(n, Der):
            if not isinstance(n.a, Leaf):
                raise ValueError
            key = n.key()
            for k, other in enumerate(ders):
                if other is n or other.key() == key:
                    break
            else:
                ders.append(n)
                k = len(ders) - 1
            code.append((nat.X_PUSH_DER, k, 0.0))
            push()
        elif isinstance(n, Leaf):
            for k, other in enumerate(leaves):
                if other.field is n.field and other.t == n.t:
                    break
            else:
                leaves.append(n)
                k = len(leaves) - 1
            code.append((nat.X_PUSH_FIELD, k, 0.0))
            push()
        elif isinstance(n, Unary):
            emit(n.a)
            code.append((nat.X_NEG, 0, 0.0))
        elif isinstance(n, Binary) and isinstance(n.a, Node) and isinstance(n.b, Node):
            emit(n.a)
            emit(n.b)
            code.append((_X_BIN[n.op], 0, 0.0))
            push(-1)
        elif isinstance(n, Binary) and isinstance(n.a, Node):
            emit(n.a)
            s = float(n.b)
            if n.op == "pow":
                if s == 1.0:
                    return
                code.append((_X_POW.get(s, nat.X_POW_IMM), 0, s))
            else:
                code.append((_X_IMM[n.op], 0, s))
        else:
            raise ValueError

    try:
        emit(node)
    except (ValueError, KeyError):
        return None
    if (len(ders) > nat.FD_MAX_TERMS or len(leaves) > nat.FD_EXPR_MAX_FIELDS or not code
            or len(code) > nat.FD_EXPR_MAX_CODE or depth[1] > nat.FD_EXPR_MAX_STACK):
        return None
    return ders, leaves, code


def top_level_plan_key(key):
    """Plan keys of fused statements are (id(field), parity, chain ids, statement key)."""
    return isinstance(key, tuple) and len(key) == 4 and isinstance(key[1], int)


def _has_nested_der(node):
    if isinstance(node, Der):
        return not isinstance(node.a, Leaf) or _has_nested_der(node.a)
    if isinstance(node, Unary):
        return _has_nested_der(node.a)
    if isinstance(node, Binary):
        return any(isinstance(s, Node) and _has_nested_der(s) for s in (node.a, node.b))
    return False


def _lift_nested_ders(node, model, t):
    """Replace the argument of every der that is not a stored level by a scratch
    leaf.  Returns (rewritten tree, [(scratch leaf, inner expression)], [the ders
    that now read a scratch leaf]); equal
    inner expressions share one scratch level.  (Ders nested deeper stay inside
    the inner expressions: they are lifted when those get their own plans.)"""
    lifted, by_key, rewritten = [], {}, []

    def walk(n):
        if isinstance(n, Der):
            if isinstance(n.a, Leaf):
                return n
            key = n.a.key()
            leaf = by_key.get(key)
            if leaf is None:
                leaf = by_key[key] = TempLeaf(_TempField(model, n.a.shape, f"scratch {len(lifted)}"), t)
                lifted.append((leaf, n.a))
            new = Der(leaf, n.shape, n.axis, n.shifts, n.weights, n.scale, n.periodic, n.lo, n.hi, n.n_src, n.n_dst,
                      n.stencil)
            new.slab = n.slab
            rewritten.append(new)
            return new
        if isinstance(n, Unary):
            a = walk(n.a)
            return n if a is n.a else Unary(n.op, a)
        if isinstance(n, Binary):
            a = walk(n.a) if isinstance(n.a, Node) else n.a
            b = walk(n.b) if isinstance(n.b, Node) else n.b
            return n if (a is n.a and b is n.b) else Binary(n.op, a, b)
        return n

    return walk(node), lifted, rewritten


def _slab_reach(stmts):
    """Largest stencil reach along the decomposed axis over the recorded statements."""
    reach = 0

    def walk(node):
        nonlocal reach
        if isinstance(node, Der):
            if node.slab:
                reach = max(reach, max(abs(s) for s in node.shifts))
            walk(node.a)
        elif isinstance(node, Unary):
            walk(node.a)
        elif isinstance(node, Binary):
            for side in (node.a, node.b):
                if isinstance(side, Node):
                    walk(side)

    for _, _, node in stmts:
        if node is not None:
            walk(node)
    return reach


def _collect_ghost_reads(node, out):
    """(id(field), level) of every stored level that ``node`` differentiates along the decomposed axis."""
    if isinstance(node, Der):
        if node.slab and isinstance(node.a, Leaf):
            out.add((id(node.a.field), node.a.t))
        _collect_ghost_reads(node.a, out)
    elif isinstance(node, Unary):
        _collect_ghost_reads(node.a, out)
    elif isinstance(node, Binary):
        for side in (node.a, node.b):
            if isinstance(side, Node):
                _collect_ghost_reads(side, out)


def _collect_new_reads(node, t_new, out):
    if isinstance(node, Leaf):
        if node.t == t_new:
            out.add(id(node.field))
    elif isinstance(node, (Unary, Der)):
        _collect_new_reads(node.a, t_new, out)
    elif isinstance(node, Binary):
        if isinstance(node.a, Node):
            _collect_new_reads(node.a, t_new, out)
        if isinstance(node.b, Node):
            _collect_new_reads(node.b, t_new, out)


# ---------------------------------------------------------------------------
# pattern matcher:  post_k(... post_1( term_0 +/- term_1 +/- ... ))
# ---------------------------------------------------------------------------

def _match_term(node):
    """der(leaf) | -der(leaf) | der(leaf)*scalar  ->  (der, pre, negate)"""
    if isinstance(node, Der) and isinstance(node.a, Leaf):
        return node, None, False
    if isinstance(node, Unary) and isinstance(node.a, Der) and isinstance(node.a.a, Leaf):
        return node.a, None, True
    if (isinstance(node, Binary) and node.op == "mul" and not isinstance(node.b, Node)
            and isinstance(node.a, Der) and isinstance(node.a.a, Leaf)):
        return node.a, float(node.b), False
    return None


def _match_sum(node):
    """Left-associated chain of +/- over terms, in evaluation order."""
    terms = []
    while True:
        single = _match_term(node)
        if single is not None:
            terms.append((single[0], False, single[1], single[2]))
            break
        if (isinstance(node, Binary) and node.op in ("add", "sub") and isinstance(node.a, Node)
                and isinstance(node.b, Node)):
            right = _match_term(node.b)
            if right is None:
                return None
            terms.append((right[0], node.op == "sub", right[1], right[2]))
            node = node.a
            continue
        return None
    terms.reverse()
    return terms


def _match_statement(node):
    """Peel scalar / coefficient-field operations off the outside of the
    expression, then require a sum of der terms.  Returns (terms, posts)."""
    posts = []
    while True:
        terms = _match_sum(node)
        if terms is not None:
            posts.reverse()
            if len(terms) == 1 and terms[0][2] is not None and not terms[0][3]:
                # a lone `der*c` is equally a post-multiplication (same rounding);
                # that reading is the kernels' canonical form
                der, subtract, pre, negate = terms[0]
                terms = [(der, subtract, None, negate)]
                posts.insert(0, (nat.POST_MUL, pre, None))
            return terms, posts
        if isinstance(node, Unary):
            posts.append((nat.POST_NEG, None, None))
            node = node.a
        elif isinstance(node, Binary) and not isinstance(node.b, Node) and node.op in _POST_OPS:
            posts.append((_POST_OPS[node.op], float(node.b), None))
            node = node.a
        elif (isinstance(node, Binary) and node.op == "mul" and isinstance(node.a, Node)
              and isinstance(node.b, Node)):
            if isinstance(node.a, Leaf) and not isinstance(node.b, Leaf):
                posts.append((nat.POST_MUL_FIELD, None, node.a))
                node = node.b
            elif isinstance(node.b, Leaf):
                posts.append((nat.POST_MUL_FIELD, None, node.b))
                node = node.a
            else:
                return None
        else:
            return None
        if len(posts) > nat.FD_MAX_POST:
            return None
