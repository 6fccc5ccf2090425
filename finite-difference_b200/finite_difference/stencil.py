"""``Stencil``: host-side weight solve (FD.py:873-962) and the ``der`` operator
(FD.py:990-1122) as a deferred node executed by the B200 kernels.

The weights are solved exactly the way the reference solves them -- Taylor
matrix on doubled coordinates, ``inv``, rounded to two decimals (FD.py:938-951)
-- because parity to 1e-12 is impossible otherwise.
"""

from __future__ import annotations

import math

import numpy as np

from ._graph import Der, FieldInstant

_AXIS_TYPES = ["cell", "edge"]


def _solve_weights(points, der_order):
    doubled = points * 2
    n = len(doubled)
    taylor = np.power(np.asmatrix(doubled), np.asmatrix(np.arange(0, n)).transpose())
    rhs = np.zeros((n, 1))
    rhs[der_order] = math.factorial(der_order) * 2 ** der_order  # 2^der: half-step units
    return np.around(np.linalg.inv(taylor) * rhs, 2).transpose().tolist()[0]


class Stencil:

    def __init__(self, sample_points: list, der_order: int, axis_type: str = "cell", der_axis_type: str = "cell"):
        """
        creates a finite difference stencil at the given coordinates and for the given derivative order

        sample_points: the coordinates for creating the stencil
        der_order: the order of the derivative
        axis_type / der_axis_type: "cell" or "edge" -- where the samples are taken / where the derivative lands.
        Same types need integer sample points, different types need half values.
        """
        assert type(sample_points) == list, \
            "points must be a list of locations to sample from for computing the derivative"
        assert type(der_order) == int and der_order > 0, "derivative order must be a positive integer"
        assert len(set(sample_points)) == len(sample_points), \
            "the coordinates of the points to sample must all be unique values"
        assert axis_type in _AXIS_TYPES, f"axis_type must be one of {_AXIS_TYPES}"
        assert der_axis_type in _AXIS_TYPES, f"der_axis_type must be one of {_AXIS_TYPES}"
        if len(sample_points) <= der_order:
            raise Exception("there must be at least one more point than the derivative order")

        self.from_edge = axis_type == "edge"
        self.to_edge = der_axis_type == "edge"
        self.points = np.array(sample_points)

        if np.all(self.points % 1 == 0):
            self.intermediate = False
        elif np.all((self.points + 0.5) % 1 == 0):
            self.intermediate = True
        else:
            raise Exception("points need to be all integers or all half values")

        # FD.py:922-933 (the cell->edge check is unreachable in the reference, FD.py:925; kept so)
        if self.from_edge and not self.to_edge and not self.intermediate:
            raise Exception(
                "going from edges to cells requires stencil with half values in difference approximation")
        if self.from_edge and self.to_edge and self.intermediate:
            raise Exception(
                "going from edges to edges requires stencil with integer values in difference approximation")
        if not self.from_edge and not self.to_edge and self.intermediate:
            raise Exception(
                "going from cells to cells requires stencil with integer values in difference approximation")

        self.weights = _solve_weights(self.points, der_order)
        self.der_order = der_order
        self._der_memo = {}
        self.to_text()

    def to_text(self):
        """
        prints the equation for the finite difference approximation
        """
        terms = "+".join(f"{w}f(x+{p}h)" for w, p in zip(self.weights, self.points))
        equation = f"f{chr(39) * self.der_order} = [{terms}] / [h^{self.der_order}]"
        for old, new in (("+-", "-"), (".0", ""), ("1f", "f"), ("1h", "h"), (")+", ") + "), (")-", ") - ")):
            equation = equation.replace(old, new)  # FD.py:981-986
        print(f"Finite approximation: {equation}\n")

    def der(stencil, f: FieldInstant, der_axis: str):
        """
        returns the derivative of f at the current time using the stencil
        performs derivative operation along der_axis
        """
        assert isinstance(f, FieldInstant), \
            "Can only perform derivative on a field at a moment, use prev or new properties of the field"
        assert isinstance(stencil, Stencil), "stencil must be a Stencil object"
        assert type(der_axis) == str, "der_axis must be a string of the axis name"
        model = f.model
        # the same call comes back every time step: everything below depends only on
        # (model, axis, where the operand lives, its shape), so a call that passed
        # all checks once is answered from the stencil's memo
        memo_key = (model._token, der_axis, tuple(f.edge_axes), f._node.shape)
        hit = stencil._der_memo.get(memo_key)
        if hit is not None:
            return FieldInstant(model, list(hit[1]), f.name, hit[0].with_operand(f._node))
        assert model.time_axis is not None, "cannot take derivative of a field with no time axis"
        space = model.axes_names
        space.remove(model.time_axis)
        assert der_axis in space, f"der_axis must be a spatial axis, '{der_axis}' is not"
        axis_n = space.index(der_axis)

        on_edges = der_axis in f.edge_axes
        if on_edges and not stencil.from_edge:
            raise Exception(
                f"Stencil takes a field with cells, but the field has edges along the {der_axis} axis")
        elif not on_edges and stencil.from_edge:
            raise Exception(
                f"Stencil takes a field with edges, but the field has cells along the {der_axis} axis")

        # FD.py:1031-1044: staggered stencils move by half a cell; truncation toward zero
        if on_edges and not stencil.to_edge:
            shifts, grow = (stencil.points + 1 / 2).astype(int), -1
        elif not on_edges and stencil.to_edge:
            shifts, grow = (stencil.points - 1 / 2).astype(int), 1
        else:
            shifts, grow = stencil.points.astype(int), 0

        src_shape = f._node.shape
        n_src = src_shape[axis_n]
        periodic = der_axis in model.periodic
        slab = model._slab
        along_slab = slab is not None and der_axis == slab.axis
        if along_slab:
            # ranges are worked out on the GLOBAL axis; the engine maps them to the slab
            n_src = slab.n + (1 if (on_edges and not periodic) else 0)
        if periodic:
            n_dst, lo, hi = n_src, 0, n_src
        else:
            n_dst = n_src + grow
            guard = 1 if (on_edges and not stencil.to_edge) else 0
            lo_shift = min(shifts[0], 0)           # first / last in USER order (FD.py:1073-1074)
            hi_shift = max(shifts[-1], guard)
            span = n_src + lo_shift - hi_shift
            lo, hi = -lo_shift, -lo_shift + span
            if span < 0:
                raise ValueError("negative dimensions are not allowed")
            for s in shifts:
                # numpy slicing rules: unsorted points give slices of unequal length
                a, b = int(s - lo_shift), int(s - lo_shift + span)
                start, stop, _ = slice(a, b).indices(n_src)
                if max(0, stop - start) != span:
                    raise ValueError(
                        f"operands could not be broadcast together with shapes ({span},) ({max(0, stop - start)},)")
                if span and start != a:
                    raise NotImplementedError(
                        "stencil sample points must be sorted ascending on a fixed axis")
        out_shape = list(src_shape)
        out_shape[axis_n] = n_dst
        if along_slab:
            lands_on_edges = stencil.to_edge if stencil.from_edge != stencil.to_edge else on_edges
            out_shape[axis_n] = slab.owned(lands_on_edges and not periodic)

        dx = dict(zip(model.axes_names, model.axes_step_size))[der_axis]
        scale = dx ** stencil.der_order            # may be np.int64, like the reference

        edges = list(f.edge_axes)
        if stencil.from_edge and not stencil.to_edge:
            edges.remove(der_axis)
        elif not stencil.from_edge and stencil.to_edge:
            edges += der_axis  # FD.py:1120 (`list += str`: one entry per character, kept for parity)

        node = Der(f._node, out_shape, axis_n, shifts, stencil.weights, scale, periodic, int(lo), int(hi),
                   int(n_src), int(n_dst), stencil)
        node.slab = along_slab
        stencil._der_memo[memo_key] = (node, tuple(edges))
        out = FieldInstant(model, edges, f.name, node)
        return out
