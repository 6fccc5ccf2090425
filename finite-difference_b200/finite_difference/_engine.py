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
        self.size = int(np.prod(field._slice_shape)) if field._slice_shape else 1
        self.constant = isinstance(field, ConstantField)
        self.integrated = field.dot is not None
        depth = 2 if self.integrated else 1
        self.ring = [engine.dev.empty(self.size, engine.dtype) for _ in range(depth)]
        self.held = [-1] * depth          # time index each slot holds
        self.saved = set()                # time indices already in the host history
        self.faces = None                 # {(dim, side): (tensor, per_time_elems or 0)}
        self.counts = engine.dev.zeros_i64(engine.n_t) if self.integrated else None
        self.counts_checked = engine.model.time_step
        s = len(field._slice_shape)
        self.dims3 = (1,) * (3 - s) + tuple(field._slice_shape) if s <= 3 else None
        if self.constant:
            self.load_level(0)
        elif self.integrated:
            self.load_level(engine.model.time_step)

    def slot_index(self, t):
        return t % len(self.ring)

    def load_level(self, t):
        """Host history -> device (initial condition, constant coefficient)."""
        eng = self.engine
        level = self.field._hist.direct_level(t)
        if level is None:
            level = self.field._hist.read({eng.model.time_axis: t})
        k = self.slot_index(t)
        eng.dev.upload(level, eng.dtype, out=self.ring[k])
        self.held[k] = t
        self.saved.add(t)
        if self.counts is not None:      # NaN count of the level, on the device (FD.py:528)
            self.counts[t] = 0
            nat.check(eng._fn["fd_count_nan"](ptr(self.ring[k]), self.size, self.counts.data_ptr() + 8 * t,
                                              eng.stream))

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
        return self.ring[k]

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
                full = np.moveaxis(full, free.index(model.time_axis), 0)
                per = int(np.prod(full.shape[1:])) if full.ndim > 1 else 1
                self.faces[(dim, side)] = (eng.dev.upload(full, eng.dtype), per)
            else:
                face = f._hist.read({axis: idx, model.time_axis: t_from})
                self.faces[(dim, side)] = (eng.dev.upload(face, eng.dtype), 0)

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


class Engine:
    def __init__(self, model):
        self.model = model
        self.dev = Device.get()
        self.dtype = model.dtype
        self.n_t = len(model.axes[model.time_axis])
        self.flags = self.dev.zeros_i64(nat.FD_NUM_FLAGS)
        self.states = {}
        self.exact_div = os.environ.get("FD_B200_FAST_SCALE", "0") != "1"
        self.fuse = os.environ.get("FD_B200_NO_FUSE", "0") != "1"
        points = max([int(np.prod(f._slice_shape)) for f in model.fields] or [1])
        default_poll = 1 if points < (1 << 20) else 64
        self.poll_every = int(os.environ.get("FD_B200_POLL_EVERY", default_poll))
        self.last_poll = model.time_step
        self.pending_copies = []          # [(state, slot, t, host tensor, event)]
        self.stats = {"fused": 0, "unfused_statements": 0}
        self._fn = {name: nat.fn(name, self.dtype) for name in
                    ("fd_der", "fd_ewise", "fd_merge_store", "fd_fill", "fd_face_store", "fd_count_nan",
                     "fd_fused_step")}

    # ------------------------------------------------------------------ utils
    def state(self, field):
        st = self.states.get(id(field))
        if st is None:
            st = self.states[id(field)] = FieldState(self, field)
        return st

    @property
    def stream(self):
        return self.dev.stream_ptr

    def wants_level(self, t):
        h = self.model.history
        if t == self.n_t - 1 or h == "all":
            return True
        if h == "last":
            return False
        return t % h == 0

    # --------------------------------------------------------------- execute
    def execute(self, stmts):
        t = self.model.time_step
        for f in self.model.fields:          # initial levels go up before the first launch
            self.state(f)
        reads_new = set()
        for _, _, node in stmts:
            if node is not None:
                _collect_new_reads(node, t + 1, reads_new)
        memo = {}
        consumed = set()
        for i, (kind, field, node) in enumerate(stmts):
            if i in consumed:
                continue
            if kind == "assign":
                chain = self._find_chain(stmts, i, field)
                if self.fuse and chain and self._try_fused(field, node, chain, t, reads_new):
                    consumed.update(j for j, _ in chain)
                    self.stats["fused"] += 1
                    continue
                self._generic_assign(field, node, t, memo)
            else:
                self._generic_integrate(field, t)
            self.stats["unfused_statements"] += 1
        for f in self.model.fields:
            st = self.state(f)
            if not st.constant and self.wants_level(t + 1):
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
    def _try_fused(self, dot_field, node, chain, t, reads_new):
        dot_state = self.state(dot_field)
        dims3 = dot_state.dims3
        if dims3 is None or dot_field._bc_locations:
            return False
        matched = _match_statement(node)
        if matched is None:
            return False
        terms, posts = matched
        if len(terms) > nat.FD_MAX_TERMS or len(posts) > nat.FD_MAX_POST:
            return False
        levels = [self.state(p) for _, p in chain]
        if any(ls.field._slice_shape != dot_field._slice_shape for ls in levels):
            return False
        if node.shape != tuple(dot_field._slice_shape):
            return False
        pad = 3 - len(dot_field._slice_shape)

        d = nat.Fused()
        for a in range(3):
            d.n[a] = dims3[a]
        d.nterms = len(terms)
        for k, (der, subtract, pre, negate) in enumerate(terms):
            leaf = der.a
            src_state = self.state(leaf.field)
            try:
                src = src_state.level(leaf.t)
            except RuntimeError:
                return False
            if max(abs(s) for s in der.shifts) >= der.n_src and der.periodic:
                return False
            tm = d.term[k]
            tm.src = src.data_ptr()
            tm.axis = der.axis + pad
            tm.periodic = 1 if der.periodic else 0
            tm.lo, tm.hi, tm.n_src = der.lo, der.hi, der.n_src
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
                    return False
                try:
                    d.post[k].coef = self.state(leaf.field).level(leaf.t).data_ptr()
                except RuntimeError:
                    return False
        d.depth = len(levels)
        d.dt = float(self.model.dt)
        want_dot = id(dot_field) in reads_new or self.wants_level(t + 1)
        d.dot_out = dot_state.claim(t + 1).data_ptr() if want_dot else 0
        if not want_dot:
            dot_state.held[0] = -1
        for l, ls in enumerate(levels):
            d.prev[l] = ls.level(t).data_ptr()
            d.out[l] = ls.claim(t + 1).data_ptr()
            d.nan_count[l] = ls.counts.data_ptr() + 8 * (t + 1)
            for (dim, side), _ in ls.face_items():
                d.face[l][dim + pad][side] = ls.face_ptr(dim, side, t + 1)
        d.flags = self.flags.data_ptr()
        nat.check(self._fn["fd_fused_step"](C.byref(d), self.stream))
        return True

    # --------------------------------------------------------- unfused path
    def _eval(self, node, memo):
        """Evaluate an expression node with one launch per reference operation."""
        got = memo.get(id(node))
        if got is not None:
            return got
        n = int(np.prod(node.shape)) if node.shape else 1
        if isinstance(node, Leaf):
            out = self.state(node.field).level(node.t)
        elif isinstance(node, Const):
            out = self.dev.upload(node.array, self.dtype)
        elif isinstance(node, Unary):
            a = self._eval(node.a, memo)
            out = self.dev.empty(n, self.dtype)
            nat.check(self._fn["fd_ewise"](nat.OP_NEG, ptr(out), ptr(a), 0, 0.0, n, self.stream))
        elif isinstance(node, Binary):
            out = self._eval_binary(node, n, memo)
        elif isinstance(node, Der):
            a = self._eval(node.a, memo)
            shape = node.a.shape
            outer = int(np.prod(shape[:node.axis])) if node.axis else 1
            inner = int(np.prod(shape[node.axis + 1:])) if node.axis + 1 < len(shape) else 1
            st = nat.make_stencil(node.shifts, node.weights, node.scale, self.exact_div)
            out = self.dev.empty(n, self.dtype)
            nat.check(self._fn["fd_der"](ptr(out), ptr(a), outer, node.n_src, node.n_dst, inner, C.byref(st),
                                          1 if node.periodic else 0, node.lo, node.hi, ptr(self.flags),
                                          self.stream))
        else:
            raise TypeError(f"unknown expression node {type(node).__name__}")
        memo[id(node)] = out
        return out

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
        nat.check(self._fn["fd_fill"](ptr(slot), float("nan"), st.size, self.stream))
        shape = st.field._slice_shape
        for (dim, side), _ in st.face_items():
            outer = int(np.prod(shape[:dim])) if dim else 1
            inner = int(np.prod(shape[dim + 1:])) if dim + 1 < len(shape) else 1
            index = 0 if side == 0 else shape[dim] - 1
            nat.check(self._fn["fd_face_store"](ptr(slot), st.face_ptr(dim, side, t), outer, shape[dim], inner,
                                                 index, self.stream))
        return slot

    def _generic_assign(self, field, node, t, memo):
        if node.shape != tuple(field._slice_shape):
            raise ValueError(f"could not broadcast input array from shape {node.shape} into shape {tuple(field._slice_shape)}")
        st = self.state(field)
        fresh = self._eval(node, memo)
        slot = self._prefilled(st, t + 1)
        nat.check(self._fn["fd_merge_store"](ptr(slot), ptr(fresh), st.size, 0, ptr(self.flags), self.stream))

    def _generic_integrate(self, field, t):
        st, dot = self.state(field), self.state(field.dot)
        fresh = self.dev.empty(st.size, self.dtype)
        nat.check(self._fn["fd_ewise"](nat.OP_EULER, ptr(fresh), ptr(st.level(t)), ptr(dot.level(t + 1)),
                                        float(self.model.dt), st.size, self.stream))
        slot = self._prefilled(st, t + 1)
        nat.check(self._fn["fd_merge_store"](ptr(slot), ptr(fresh), st.size, st.counts.data_ptr() + 8 * (t + 1),
                                              ptr(self.flags), self.stream))

    # ------------------------------------------------------ host interaction
    def evaluate_to_host(self, node):
        out = self._eval(node, {})
        arr = self.dev.download(out, node.shape)
        self.poll(final=True)
        return np.array(arr, dtype=np.float64)

    def snapshot(self, st, t):
        """Device level -> host history."""
        k = st.slot_index(t)
        if t in st.saved or st.held[k] != t:
            return
        arr = self.dev.download(st.ring[k], st.field._slice_shape)
        st.field._hist.write_level(t, np.array(arr, dtype=np.float64))
        st.saved.add(t)

    def wait_for_copies(self, st, slot):
        return

    def sync_history(self, field):
        """Bring ``Field.data`` up to date: outstanding copies plus the level the
        model currently stands on."""
        st = self.states.get(id(field))
        if st is None or st.constant:
            return
        self.snapshot(st, self.model.time_step)
        self.poll(final=True)

    def host_changed(self, field):
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
        flags = self.flags.cpu().numpy()    # synchronises the compute stream
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
            counts = st.counts[st.counts_checked:now + 1].cpu().numpy()
            first = st.counts_checked
            st.counts_checked = now
            for k in range(1, len(counts)):
                # FD.py:535: only from the second step on, and only once per model
                if first + k > 1 and counts[k] > counts[k - 1]:
                    print(LOSING_MSG)
                    self.model.insufficient_bc = True
                    return


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
            # a lone `der*scalar` term is equally a post-multiplication; either
            # reading rounds identically
            posts.reverse()
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
