"""``Model`` (FD.py:14-258): axes, time cursor and the place where the recorded
statements of a time step are handed to the GPU engine.

Additive options (defaults reproduce the reference):
  dtype     np.float64 (reference) or np.float32
  history   "all" (reference: every level of every field comes back to
            ``Field.data``), an integer K (every K-th level plus the last
            one), or "last" (nothing is streamed; ``Field.data`` fetches the level
            the model currently stands on; ``.dot`` levels are not kept)
"""

from __future__ import annotations

import numbers
import os

import numpy as np


class Model:

    def __init__(self, axes: dict, time_axis: str, periodic: tuple = (), *, dtype=None, history=None):
        """
        creates the domain which will be used for the simulation

        axes: the keys are the name of the axis, and the values are an evenly \
            spaced array of values for the given axis
        time_axis: the axis name that represents time
        periodic: tuple of axes names that are periodic
        """
        assert type(axes) == dict, "axes must be a dictionary of axes and corresponding values"
        assert type(time_axis) == str, "time_axis must be a string"
        if type(periodic) == list:
            periodic = tuple(periodic)
        assert type(periodic) == tuple, "periodic must be a list or tuple"

        for axis_name, axis_vals in axes.items():
            msg = _axis_problem(axis_name, axis_vals)
            if msg is not None:
                raise Exception(f"error with axis {axis_name}: {msg}")
            axes[axis_name] = np.array(axis_vals)  # the caller's dict is updated, as in FD.py:66

        assert type(time_axis) == str, "time axis must be a string"
        assert time_axis in axes, f"time axis must be one of the axes names, \
{time_axis} is not"
        assert time_axis not in periodic, "time axis can't be periodic"
        non_axes_periodic = set(periodic) - set(axes.keys())
        assert non_axes_periodic == set(), f"periodic axes must be one of the axes names, \
{non_axes_periodic} is/are not"

        self.axes = axes
        # derived axis facts are fixed after construction; the reference
        # recomputes them on every access (FD.py:243-258), which costs O(axis
        # length) Python work per call and would dominate large-grid steps
        self._names = list(axes.keys())
        self._lengths = [len(v) for v in axes.values()]
        self._steps = [v[1] - v[0] for v in axes.values()]
        self._token = object()          # identity of this model in memo keys (ids can be reused, this cannot)
        self._t_pos = self._names.index(time_axis)
        self.periodic = periodic
        self.time_step = 0
        self.time_axis = time_axis
        self.fields = []
        self.insufficient_bc = False

        self.dtype = np.dtype(dtype if dtype is not None else os.environ.get("FD_B200_DTYPE", "float64"))
        assert self.dtype in (np.dtype(np.float64), np.dtype(np.float32)), "dtype must be float64 or float32"
        history = history if history is not None else os.environ.get("FD_B200_HISTORY", "all")
        if isinstance(history, str) and history.isdigit():
            history = int(history)
        assert history in ("all", "last") or (isinstance(history, int) and history >= 1), \
            "history must be 'all', 'last' or a positive integer"
        self.history = history
        self._pending = []       # statements recorded in the current step
        self._engine = None      # created at the first executed statement
        from ._dist import make_slab
        self._slab = make_slab(self)   # None unless this process is one rank of several

    # ---- reference API -----------------------------------------------------
    def check_IC(self):
        """
        checks if all fields with time derivatives have had their initial conditions set
        """
        for field in self.fields:
            if field.dot is not None and not field.IC_set:
                raise Exception(f"field {field.name} has not had its initial conditions set")
        print("All initial conditions set!\n")

    def increment_time(self):
        """
        must be called for each iteration of the simulation
        increments time step and resets all field's updated property to False
        """
        from .field import ConstantField
        for field in self.fields:
            if isinstance(field, ConstantField):
                continue
            assert field.updated, f"all fields must be updated in the iteration, {field.name} has not been"
        self._flush()
        for field in self.fields:
            if not isinstance(field, ConstantField):
                field.updated = False
        self.time_step += 1
        if self._engine is not None:
            self._engine.step_closed()

    @property
    def finished(self):
        """
        true only when the simulation has finished running
        """
        n_steps = len(self.axes[self.time_axis])
        if self.time_step >= n_steps:
            raise ValueError(
                "the current timestep reached the number of timestips -- this shouldn't happen")
        done = self.time_step + 1 == n_steps
        if done and self._engine is not None:
            self._engine.poll(final=True)
        return done

    @property
    def dt(self):
        assert self.time_axis != None, "no time axis"
        return dict(zip(self.axes_names, self.axes_step_size))[self.time_axis]

    @property
    def n_time_steps(self):
        assert self.time_axis != None, "no time axis"
        return dict(zip(self.axes_names, self.axes_lengths))[self.time_axis]

    @property
    def axes_lengths(self):
        return list(self._lengths)

    @property
    def axes_names(self):
        return list(self._names)

    @property
    def axes_values(self):
        return list(self.axes.values())

    @property
    def axes_step_size(self):
        return list(self._steps)

    def interact(self):
        """
        generates interactive visual of the fields over time (jupyter + matplotlib + ipywidgets)
        """
        from ._interact import interact
        interact(self)

    # ---- engine hooks --------------------------------------------------------
    def _get_engine(self):
        if self._engine is None:
            from ._engine import Engine
            self._engine = Engine(self)
        return self._engine

    def _record_assign(self, field, instant):
        self._pending.append(("assign", field, instant._node))

    def _record_integrate(self, field):
        self._pending.append(("integrate", field, None))

    def _flush(self):
        if self._pending:
            stmts, self._pending = self._pending, []
            self._get_engine().execute(stmts)

    def _materialise(self, node):
        """Values of an expression node as a host array (``FieldInstant.data``)."""
        from ._graph import Leaf
        if isinstance(node, Leaf) and self._engine is None and not self._pending:
            # nothing has run yet: the host history is authoritative
            return node.field._hist.read({self.time_axis: node.t})
        self._flush()
        return self._get_engine().evaluate_to_host(node)

    def _sync_history(self, field):
        if self._engine is not None:
            self._flush()
            self._engine.sync_history(field)

    def _host_changed(self, field):
        if self._engine is not None:
            self._engine.host_changed(field)

    def synchronize(self):
        """Wait for all queued GPU work and raise any deferred error (additive)."""
        self._flush()
        if self._engine is not None:
            self._engine.poll(final=True)

    def run(self, body, steps=None):
        """Additive: ``steps`` time steps (default: all that remain) of a loop whose
        body is the same every step,

            m.run(body)      instead of      while not m.finished: body(); m.increment_time()

        ``body`` is called for the first two steps only (once per parity of the
        two-level device rings); the statements it recorded are then replayed with
        their time levels moved on, without going through the Python operators
        again -- the host cost of a step drops from ~45 us to ~15 us, which is what
        bounds small grids and strong scaling.  Everything a step does otherwise
        (history policy, polling of the deferred conditions, boundary values that
        depend on time) happens as in the written-out loop.  A body whose
        statements depend on the step number through Python values (a coefficient
        computed from ``m.time_step``, an ``if``) must use the written-out loop."""
        from ._graph import stored_level_leaves
        n_t = len(self.axes[self.time_axis])
        remaining = n_t - 1 - self.time_step if steps is None else int(steps)
        assert 0 <= remaining <= n_t - 1 - self.time_step, "not that many time steps left"
        recorded = {}
        while remaining > 0:
            parity = self.time_step & 1
            entry = recorded.get(parity)
            if entry is None:
                # record this parity's statements while executing them the usual way
                body()
                stmts = list(self._pending)
                self.increment_time()
                recorded[parity] = [stmts, self.time_step - 1, self._engine.last_keys, stored_level_leaves(stmts)]
            else:
                stmts, at, keys, leaves = entry
                by = self.time_step - at
                for leaf in leaves:
                    leaf.t += by
                entry[1] = self.time_step
                self._engine.execute(stmts, keys)
                self.time_step += 1
                self._engine.step_closed()
            remaining -= 1

    def close(self):
        """Release the model's device state (additive).  ``Field.data`` is brought
        up to date first, so results stay readable; the model cannot be advanced
        afterwards.  On several GPUs every rank closes its model at the same point
        of the script; buffers that neighbouring GPUs have mapped are kept for the
        next model of the same geometry instead of being freed."""
        if self._engine is None:
            return
        self.synchronize()
        if self.history != "last":
            for field in self.fields:
                self._engine.sync_history(field)
        self._engine.close()
        # earlier closed models are cycles of Python objects that still own device
        # and pinned memory until the cycle collector has run
        import gc
        gc.collect()


def _axis_problem(axis_name, vals):
    """First problem with one axis, or None (FD.py:36-58)."""
    if type(axis_name) != str:
        return "axes keys must be strings, the name of the axis"
    try:
        iter(vals)
    except TypeError:
        return "axes must be iterable"
    if not np.all([isinstance(val, numbers.Number) for val in vals]):
        return "all values of axis must be numbers"
    if len(vals) < 2:
        return "must have more than one value in axis"
    diffs = np.diff(vals)
    tolerance = 10 ** -9
    if not np.all(np.abs(diffs - diffs[0]) < tolerance):
        return "all values of axes must be evenly spaced"
    if diffs[0] < tolerance:
        return "axis values cannot be repeated"
    return None
