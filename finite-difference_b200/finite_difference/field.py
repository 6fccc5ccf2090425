"""``Field`` and ``ConstantField`` (FD.py:260-772) on top of a device ring
buffer.

The public surface is the reference's: ``dot``, ``prev``, ``new``, ``set_IC``,
``set_BC``, ``assign_update``, ``time_integrate_update``, ``data``, ``imshow``,
``plot`` and the attributes ``name, model, edge_axes, updated, IC_set``.  The
bodies differ: updates are *recorded* and executed on the GPU when the step is
closed, and ``data`` is assembled from the host history on demand.
"""

from __future__ import annotations

import numpy as np

from . import _expr
from ._graph import FieldInstant, Leaf
from ._history import HostHistory


class Field:
    _is_constant = False

    def __init__(self, model, name: str, edge_axes: tuple = (), n_time_ders: int = 0):
        """
        Fields are scalar fields which are updated during the simulation

        model: a Model object which the field is part of, all fields should share the same model
        name: field name, used for labeling plots
        edge_axes: the axes along which the values of the field are at the edges, between cells
        n_time_ders: number of time derivatives of the field which are used for modelling,
        use the dot property to get the time derivative of a field
        """
        from .model import Model
        assert isinstance(model, Model), "model must be a Model object"
        assert type(name) == str, "field name must be a string"
        assert type(edge_axes) in [str, list, tuple], "edge axes must be a string, list or tuple"
        assert set(edge_axes) - set(model.axes_names) == set(), "edge_axes must be axes in the model"
        assert model.time_axis not in edge_axes, "edge_axes cannot include the time axis"
        assert type(n_time_ders) == int and n_time_ders >= 0, \
            "n_time_ders must be an integer greater than equal to 0"

        edge_axes = tuple(edge_axes)
        self.name = name
        # time derivatives are fields of their own, created first (FD.py:293-298)
        self.dot = Field(model, f"{self.name} dot", edge_axes, n_time_ders - 1) if n_time_ders > 0 else None
        self._parent = None
        if self.dot is not None:
            self.dot._parent = self
        model.fields.append(self)

        lengths = dict(zip(model.axes_names, model.axes_lengths))
        for axis in edge_axes:
            if axis not in model.periodic:
                lengths[axis] += 1      # both end edges are stored (FD.py:300-304)

        self.model = model
        self.edge_axes = edge_axes
        self.updated = False
        self.IC_set = False
        self._global_shape = tuple(lengths.values())
        slab = model._slab
        if slab is not None:
            # this rank holds one slab of the slowest spatial axis (see _dist.py)
            self._slab_extra = slab.axis in edge_axes and slab.axis not in model.periodic
            lengths[slab.axis] = slab.owned(self._slab_extra)
        self._shape = tuple(lengths.values())
        self._slice_shape = tuple(n for a, n in lengths.items() if a != model.time_axis)
        self._hist = HostHistory(model.axes_names, self._shape, model.time_axis)
        self._bc_locations = []         # [(axis name, index)] registered by set_BC
        self._state = None              # device-side state, created by the engine

    # ---- history ---------------------------------------------------------
    @property
    def data(self):
        """Dense ``[axes in dict order, incl. time]`` float64 array, NaN where
        unknown (FD.py:310).  Levels the history policy did not keep read as
        the reference's pre-populated content (NaN plus boundary values)."""
        self.model._sync_history(self)
        slab = self.model._slab
        if slab is not None:
            return slab.gather_arrays(self._hist.materialise(), self.model._names.index(slab.axis))
        return self._hist.materialise()

    @data.setter
    def data(self, value):
        value = np.asarray(value, dtype=np.float64)
        assert value.shape == self._shape, f"data must have shape {self._shape}"
        self._hist.materialise()
        self._hist.dense = np.array(value)
        self.model._host_changed(self)

    def data_level(self, t):
        """Additive to the reference API: one time level as a host array, without
        assembling the dense history (``Field.data`` of a 512^3 x 100-step run
        would need 100 GiB).  Levels the history policy skipped read as NaN."""
        t = int(t) + (self._shape[self.model.axes_names.index(self.model.time_axis)] if t < 0 else 0)
        level = self.local_level(t)
        slab = self.model._slab
        if slab is not None:
            space = [a for a in self.model._names if a != self.model.time_axis]
            return slab.gather_arrays(level, space.index(slab.axis))
        return level

    def local_level(self, t, out=None):
        """Additive: like ``data_level`` but only this rank's slab when the model
        is decomposed over several GPUs (no communication).  ``out``: a host
        array of the slab's shape and the model's dtype to receive the level the
        model currently stands on straight from the device (pinned memory makes
        this a single DMA transfer; nothing is allocated)."""
        t = int(t) + (self._shape[self.model.axes_names.index(self.model.time_axis)] if t < 0 else 0)
        if out is not None and self.model._engine is not None and t == self.model.time_step:
            self.model._flush()
            if self.model._engine.fetch_into(self, t, out):
                return out
        self.model._sync_history(self)
        level = self._hist.direct_level(t)
        return level if level is not None else self._hist.read({self.model.time_axis: t})

    # ---- time levels -----------------------------------------------------
    @property
    def new(self):
        """
        Returns the section of the field at the upcoming time
        Can only be used after the field has been updated
        """
        assert self.updated, "field has not yet been updated, use prev property to get current values"
        return self._instant(self.model.time_step + 1)

    @property
    def prev(self):
        """
        Returns the section of the field at the current time
        Can only be used before the field has been updated
        """
        assert not self.updated, "field has already been updated, use new property to get values at the next timestep\n\
            alternatively, you may have forgot to call the increment_time method at the end of the iteration "
        assert self.dot is not None, "cannot access prev for a field without a time derivative since no initial conditions are set,\n\
instead first use the assign_update method and then use the new property"
        return self._instant(self.model.time_step)

    def _instant(self, t):
        model = self.model
        # (_check_index for the time axis alone, FD.py:596-604)
        assert model._lengths[model._t_pos] > t, \
            f"{model.time_axis} goes out of bounds, it has length {model._lengths[model._t_pos]}"
        return FieldInstant(model, self.edge_axes, self.name, Leaf(self, t))

    def _check_index(self, idxs):
        # FD.py:596-604
        assert type(idxs) == dict, "indexing must be a dictionary"
        lengths = dict(zip(self.model.axes_names, self.model.axes_lengths))
        for axis in idxs:
            assert type(axis) == str, f"{axis} is not a string"
            assert axis in self.model.axes_names, f"{axis} is not an axis in the model"
            assert lengths[axis] > idxs[axis], f"{axis} goes out of bounds, it has length {lengths[axis]}"

    # ---- initial and boundary conditions ----------------------------------
    def set_IC(self, expression: str):
        """
        sets the values of the field at t=0
        only can be performed on fields with time derivatives

        expression: the mathematical expression for the field, as a function of the spatial axes
        """
        assert self.dot != None, "field must have a time derivative to set initial conditions"
        time_axis = self.model.time_axis
        assert time_axis != None, "the model must have a time axis to set the initial conditions"
        self._set_expression(expression, {time_axis: 0})
        self.IC_set = True

    def set_IC_values(self, values):
        """Additive to the reference API: initial condition from an array (the
        reference only takes strings, so its users write ``Field.data[t=0]``
        directly, which needs the whole dense history in host memory).  The
        array is used as is -- pass a view of pinned memory for fast upload."""
        assert self.dot != None, "field must have a time derivative to set initial conditions"
        values = np.asarray(values)
        slab = self.model._slab
        if slab is not None and values.shape != self._slice_shape:
            # a global array: keep this rank's slab (the slab axis is the first spatial axis)
            values = values[slab.lo:slab.lo + self._slice_shape[0]]
        assert values.shape == self._slice_shape, f"initial values must have shape {self._slice_shape}"
        self._hist.write({self.model.time_axis: 0}, values, copy=False)
        self.model._host_changed(self)
        self.IC_set = True

    def set_BC(self, expression: str, axis: str, side: str):
        """
        sets the values of the field at boundaries of the domain, for every time level

        expression: the mathematical expression for the field, as a function of the axes
        axis: the axis which the boundary condition is applied to
        side: "start" or "end", which side the boundary condition is applied to

        the expression cannot be a function of the axis
        """
        assert axis != self.model.time_axis, "cannot use BC to set time axis, use set_IC instead"
        assert axis in self.model.axes_names, "axis must be one of the axes names"
        assert axis not in self.model.periodic, "cannot set boundary conditions to periodic axis"
        side_types = ["start", "end"]
        assert side in side_types, f"side must be one of {side_types}"
        idx = 0 if side == "start" else -1
        self._set_expression(expression, {axis: idx})
        if (axis, idx) not in self._bc_locations:
            self._bc_locations.append((axis, idx))

    def _set_expression(self, expression: str, location: dict = {}):
        """Evaluate ``expression`` on the region and merge it in (FD.py:387-476)."""
        assert type(expression) == str, "expression must be a string"
        assert type(location) == dict
        model = self.model
        slab = model._slab
        if slab is not None and slab.axis in location:
            # a face of the decomposed axis belongs to the first / last rank only
            if not (slab.first if location[slab.axis] == 0 else slab.last):
                self._check_index(location)
                return
        steps = dict(zip(model.axes_names, model.axes_step_size))
        free, coords = [], []
        for name, values in zip(model.axes_names, model.axes_values):
            if name in location:
                continue
            if name in self.edge_axes:
                # edge i sits half a cell to the left of cell i (FD.py:428-441)
                dx = steps[name]
                values = [v - dx / 2 for v in values]
                if name not in model.periodic:
                    values.append(values[-1] + dx)
            if slab is not None and name == slab.axis:
                values = values[slab.lo:slab.lo + self._shape[model._names.index(name)]]
            free.append(name)
            coords.append(values)
        values = _expr.evaluate(expression, model.axes_names, free, coords)
        self._check_index(location)
        self._merge_host(values, location, allow_override=True, label=self.name)

    def _merge_host(self, values, location, allow_override, label):
        """Host-side ``__set_data`` (FD.py:541-576), used for ICs and BCs."""
        hist = self._hist
        if hist.dense is None and not hist.log and not hist.levels:
            overriding = False      # nothing stored yet: skip materialising the region
        else:
            existing = hist.read(location)
            known = ~np.isnan(np.broadcast_to(values, existing.shape))
            overriding = not np.all(np.isnan(existing[known]))
        if allow_override and overriding:
            print(f"Boundary conditions and initial conditions may be in conflict for field {label}\n\
Conflicting values override and become equal to whatever was assigned last\n")
        elif not allow_override and overriding:
            raise Exception("attempting to override values, this may be due to too many boundary conditions.")
        self._hist.write(location, values)
        self.model._host_changed(self)

    # ---- updates -----------------------------------------------------------
    def assign_update(self, field_slice):
        """
        assigns the values of field_slice to a field at a current time
        field can only be updated once each iteration
        """
        assert not self.updated, "field already updated, you may have forgot to call increment_time on the model"
        assert self.dot == None, "cannot use assign_update on field with time derivative, use time_integrate_update instead"
        assert isinstance(field_slice, FieldInstant), "invalid argument, instead use the 'prev' or 'new' properties of a field (or operations on them)"
        time_axis = self.model.time_axis
        # FD.py:550-553
        assert set(self.edge_axes) - {time_axis} == set(field_slice.edge_axes) - {time_axis}, \
            "field edge axes and field slice edge axes must match"
        self.model._record_assign(self, field_slice)
        self.updated = True

    def time_integrate_update(self):
        """
        updates a field using Euler integration of it's time derivative
        field can only be updated once each iteration
        """
        assert not self.updated, "field already updated"
        assert self.dot != None, "field needs to have time derivative to perform time_integrate"
        assert self.dot.updated, "time derivative needs to be updated first, can be accessed with 'dot' property"
        self.model._record_integrate(self)
        self.updated = True

    # ---- plotting (out of scope as compute: thin host views of .data) -----
    def _view(self, location):
        self._check_index(location)
        sel = tuple(location[a] if a in location else slice(None) for a in self.model.axes)
        return self.data[sel]

    def imshow(self, location={}):
        im_axes = [a for a in self.model.axes_names if a not in location]
        if len(im_axes) == 0:
            raise Exception("no remaining data once location is specified")
        elif len(im_axes) == 1:
            raise Exception("only 2-dimensional data allowed, use plot for 1-dimensional")
        elif len(im_axes) > 2:
            raise Exception("cannot suppport more than 2-dimensional data for imshow")
        plt = _pyplot()
        values = dict(zip(self.model.axes_names, self.model.axes_values))
        extent = [b for a in im_axes for b in (min(values[a]), max(values[a]))]
        plt.imshow(np.flipud(np.transpose(self._view(location))), extent=extent)  # FD.py:666
        plt.xlabel(im_axes[0])
        plt.ylabel(im_axes[1])

    def plot(self, location={}):
        plot_axes = [a for a in self.model.axes_names if a not in location]
        values = self._view(location)
        assert len(plot_axes) == 1, "only 1-dimensional data allowed for plot"
        plt = _pyplot()
        plt.plot(dict(zip(self.model.axes_names, self.model.axes_values))[plot_axes[0]], values)
        plt.xlabel(plot_axes[0])

    def __str__(self) -> str:
        return f"{len(self.model.axes_names)}-dimensional Field named {self.name}, dimension lengths: {dict(zip(self.model.axes_names,self.model.axes_lengths))}"

    # ---- arithmetic on a Field itself is an error (FD.py:702-739) ----------
    @staticmethod
    def _no_arithmetic():
        raise TypeError(
            "Cannot perform arithmetic between fields directly. First use the \
prev or new properties to get the fields at the current timestep.")

    def __neg__(self):
        Field._no_arithmetic()

    def __add__(self, other):
        Field._no_arithmetic()

    # the reference declares the remaining operators without `other`
    # (FD.py:714-739), so Python raises an arity TypeError instead
    def __sub__(self):
        Field._no_arithmetic()

    __mul__ = __truediv__ = __pow__ = __radd__ = __rsub__ = __rmul__ = __rtruediv__ = __rpow__ = __sub__


class ConstantField(Field):
    _is_constant = True

    def __init__(self, model, name: str, edge_axes: tuple = ()):
        """
        A Field which does not change with time
        """
        super().__init__(model, name, edge_axes, 0)
        self.updated = True

    @property
    def prev(self):
        raise Exception("prev property not allowed for ConstantField, use always property instead")

    @property
    def always(self):
        return self._instant(0)  # could choose any time (FD.py:759-762)

    def set_IC(self, _):
        raise Exception("set_IC not allowed for ConstantField, use set instead")

    def set_BC(self, _):
        raise Exception("set_BC not allowed for ConstantField, use set instead")

    def set(self, expression):
        self._set_expression(expression)


_graph_module = __import__(__name__.rsplit(".", 1)[0] + "._graph", fromlist=["Field"])
_graph_module.Field = Field


def _pyplot():
    try:
        import matplotlib.pyplot as plt
    except ImportError as exc:  # plotting is optional in the B200 build
        raise ImportError("matplotlib is required for imshow/plot/interact") from exc
    return plt
