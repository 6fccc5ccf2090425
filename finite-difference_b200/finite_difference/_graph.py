"""Deferred expression layer: ``FieldInstant`` (FD.py:774-869) as a DAG node.

In the reference every operator allocates a fresh numpy array.  Here the same
operators, with the same checks and exceptions raised at the same call, only
record what was asked for; the model executes the recorded statements on the
GPU when the time step is closed (``Model.increment_time``) or when a result
is actually looked at (``FieldInstant.data``).
"""

from __future__ import annotations

import numbers

import numpy as np


Field = None   # bound by field.py once the class exists (avoids a per-call import)


class Node:
    """Expression node.  ``shape`` is the spatial shape (model order, time
    axis removed); ``dev`` caches the device tensor once evaluated."""
    __slots__ = ("shape", "dev", "epoch")

    def __init__(self, shape):
        self.shape = tuple(shape)
        self.dev = None
        self.epoch = -1


class Leaf(Node):
    """One stored time level of a field (``prev``/``new``/``always``)."""
    __slots__ = ("field", "t")

    def __init__(self, field, t):
        super().__init__(field._slice_shape)
        self.field = field
        self.t = int(t)

    def key(self):
        # relative to the model's time cursor, so that the same statement has the
        # same key every step; a coefficient field (ConstantField.always) is one
        # fixed level whatever the step
        if self.field._is_constant:
            return ("leaf", id(self.field), "const")
        return ("leaf", id(self.field), self.t - self.field.model.time_step)


class Unary(Node):
    __slots__ = ("op", "a")

    def __init__(self, op, a):
        super().__init__(a.shape)
        self.op, self.a = op, a

    def key(self):
        return ("un", self.op, self.a.key())


class Binary(Node):
    """``a op b``; exactly one of a, b may be a Python float (scalar operand)."""
    __slots__ = ("op", "a", "b")

    def __init__(self, op, a, b):
        super().__init__(a.shape if isinstance(a, Node) else b.shape)
        self.op, self.a, self.b = op, a, b

    def key(self):
        ka = self.a.key() if isinstance(self.a, Node) else ("s", self.a)
        kb = self.b.key() if isinstance(self.b, Node) else ("s", self.b)
        return ("bin", self.op, ka, kb)


class Der(Node):
    """``stencil.der(a, axis)`` (FD.py:990-1122) with everything the kernels
    need resolved at record time."""
    __slots__ = ("a", "axis", "shifts", "weights", "scale", "periodic", "lo", "hi", "n_src", "n_dst", "stencil",
                 "slab")

    def __init__(self, a, shape, axis, shifts, weights, scale, periodic, lo, hi, n_src, n_dst, stencil):
        super().__init__(shape)
        self.a, self.axis = a, axis
        self.shifts = tuple(shifts.tolist()) if hasattr(shifts, "tolist") else tuple(int(s) for s in shifts)
        self.weights, self.scale = tuple(weights), scale
        self.periodic, self.lo, self.hi, self.n_src, self.n_dst = periodic, lo, hi, n_src, n_dst
        self.stencil = stencil
        self.slab = False   # True: along the axis that is decomposed over ranks

    def key(self):
        return ("der", self.axis, self.shifts, self.weights, float(self.scale), self.periodic, self.a.key())

    def with_operand(self, a):
        """The same derivative of another operand of the same shape (Stencil.der memo)."""
        n = Der.__new__(Der)
        n.shape, n.dev, n.epoch = self.shape, None, -1
        n.a, n.axis, n.shifts, n.weights, n.scale = a, self.axis, self.shifts, self.weights, self.scale
        n.periodic, n.lo, n.hi, n.n_src, n.n_dst = self.periodic, self.lo, self.hi, self.n_src, self.n_dst
        n.stencil, n.slab = self.stencil, self.slab
        return n


def stored_level_leaves(stmts):
    """The distinct leaves (stored time levels of fields that change in time) that
    the recorded statements read: Model.run replays a step's statements with these
    moved on in time instead of recording the statements anew."""
    seen, leaves = set(), []

    def walk(n):
        if id(n) in seen:
            return
        seen.add(id(n))
        if isinstance(n, Leaf):
            if not n.field._is_constant:
                leaves.append(n)
        elif isinstance(n, (Unary, Der)):
            walk(n.a)
        elif isinstance(n, Binary):
            if isinstance(n.a, Node):
                walk(n.a)
            if isinstance(n.b, Node):
                walk(n.b)

    for _, _, node in stmts:
        if node is not None:
            walk(node)
    return leaves


class FieldInstant:
    """Field at one instant of time (FD.py:774-788).  Same attributes as the
    reference -- ``model``, ``edge_axes``, ``name``, ``data`` -- but ``data``
    is computed on the device on first access."""

    def __init__(self, model, edge_axes, name, data):
        self.model = model
        self.edge_axes = edge_axes
        self.name = name
        if isinstance(data, Node):
            self._node = data
            self._data = None
        else:
            # constructed by a user from an array (not recommended by the
            # reference either): becomes a constant leaf uploaded on demand
            self._node = Const(np.asarray(data, dtype=np.float64))
            self._data = self._node.array

    @property
    def data(self):
        if self._data is None:
            self._data = self.model._materialise(self._node)
        return self._data

    @data.setter
    def data(self, value):
        self._data = value

    def __str__(self) -> str:  # FD.py:790-792
        return f"{len(self.model.axes_names)}-dimensional field at the current time,\n\
Dimension lengths: {dict(zip(self.model.axes_names,self.model.axes_lengths))}"

    # ---- arithmetic (FD.py:794-869) --------------------------------------
    @staticmethod
    def _operand(operand, models):
        cls = type(operand)
        if cls is FieldInstant:                 # the two everyday cases first
            models.append(operand.model)
            return operand._node
        if cls is float or cls is int:
            return float(operand)
        if isinstance(operand, Field):
            raise TypeError(
                "Cannot perform arithmetic between a field and a field at a given moment -- use the new or prev properties for both fields")
        if isinstance(operand, FieldInstant):
            models.append(operand.model)
            return operand._node
        elif isinstance(operand, numbers.Number):
            if isinstance(operand, numbers.Complex) and not isinstance(operand, numbers.Real):
                raise TypeError("complex scalars are not supported by the B200 kernels")
            return float(operand)
        else:
            raise TypeError(
                "can only perform arithmetic between fields or between fields and numbers")

    def _field_op(op1, op2, operation):
        if op2 is None:
            node = Unary(operation, op1._node)
        else:
            if isinstance(op2, np.ndarray):
                # `op2 == None` in the reference (FD.py:812) is ambiguous for arrays
                raise ValueError("The truth value of an array with more than one element is ambiguous. "
                                 "Use a.any() or a.all()")
            models = []
            a = FieldInstant._operand(op1, models)
            b = FieldInstant._operand(op2, models)
            both_fields = len(models) == 2
            if both_fields and models[0] != models[1]:
                raise ValueError(
                    "can only perform arithmetic operations between fields with the same model")
            elif both_fields and set(op1.edge_axes) != set(op2.edge_axes):
                raise Exception(
                    "Fields must share the same axes that are on edges vs on cells")
            node = Binary(operation, a, b)
        return FieldInstant(op1.model, op1.edge_axes, op1.name, node)

    def __neg__(self):
        return FieldInstant._field_op(self, None, "neg")

    def __add__(self, other):
        return FieldInstant._field_op(self, other, "add")

    def __sub__(self, other):
        return FieldInstant._field_op(self, other, "sub")

    def __mul__(self, other):
        return FieldInstant._field_op(self, other, "mul")

    def __truediv__(self, other):
        assert not isinstance(other, Field), "cannot divide by field"
        return FieldInstant._field_op(self, other, "truediv")

    def __pow__(self, other):
        assert not isinstance(other, Field), "cannot raise to field"
        return FieldInstant._field_op(self, other, "pow")

    def __radd__(self, other):
        return FieldInstant._field_op(self, other, "add")

    def __rsub__(self, other):
        # the reference computes self - other here (FD.py:859-860); kept for parity
        return FieldInstant._field_op(self, other, "sub")

    def __rmul__(self, other):
        return FieldInstant._field_op(self, other, "mul")

    def __rtruediv__(self, other):
        raise Exception("cannot divide by field")

    def __rpow__(self, other):
        raise Exception("cannot raise to field")


class Const(Node):
    """A user-supplied array wrapped in a FieldInstant."""
    __slots__ = ("array",)

    def __init__(self, array):
        super().__init__(array.shape)
        self.array = array

    def key(self):
        return ("const", id(self.array))
