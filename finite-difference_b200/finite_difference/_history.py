"""Host-side time history of one field (what the reference keeps in
``Field.data``, FD.py:310).

The reference allocates ``np.full(all axes incl. time, nan)`` up front and
writes initial conditions, boundary faces (for *every* time level) and every
computed step into it.  That is 537 GB per field for the 8192^2 x 1001
benchmark, so this class keeps the same information in two interchangeable
forms:

* a **write log** of axis-aligned regions with broadcastable values (a
  Dirichlet face "0" for all times is one 8-byte entry), replayed on demand
  with the reference's merge rule (FD.py:556-576: known values overwrite,
  NaN values leave what is there), plus complete time **levels** that came
  back from the device (snapshots), which replace whatever the log says;
* a **dense** ndarray with exactly the reference's shape and contents, created
  the first time the user touches ``Field.data`` (and then authoritative, so
  array writes into it before the first step are honoured).
"""

from __future__ import annotations

import numpy as np


class Entry:
    __slots__ = ("fixed", "values")

    def __init__(self, fixed, values):
        self.fixed = fixed      # {axis name: non-negative index}
        self.values = values    # broadcastable over the free axes, model order


def _merge_into(target, values):
    """target[known] = values[known] -- the allow_override merge of FD.py:574-576."""
    values = np.broadcast_to(values, target.shape)
    known = ~np.isnan(values)
    target[known] = values[known]


class HostHistory:
    def __init__(self, axes_names, shape, time_axis, dtype=np.float64):
        self.names = list(axes_names)
        self.shape = tuple(int(s) for s in shape)
        self.time_axis = time_axis
        self.dtype = np.dtype(dtype)
        self.log = []
        self.levels = {}     # time index -> complete level (spatial axes, model order)
        self.dense = None

    # -- helpers -----------------------------------------------------------
    def normalise(self, location):
        out = {}
        for axis, idx in location.items():
            n = self.shape[self.names.index(axis)]
            idx = int(idx)
            out[axis] = idx + n if idx < 0 else idx
        return out

    def free_axes(self, fixed):
        return [a for a in self.names if a not in fixed]

    def region_shape(self, fixed):
        return tuple(s for a, s in zip(self.names, self.shape) if a not in fixed)

    def _index(self, fixed):
        return tuple(fixed[a] if a in fixed else slice(None) for a in self.names)

    @property
    def nbytes(self):
        return int(np.prod(self.shape)) * 8

    # -- replay ------------------------------------------------------------
    def _replay(self, e_fixed, e_values, replace, fixed, free, out):
        """Apply one entry (restricted to the region ``fixed``) onto ``out``."""
        for a, i in e_fixed.items():      # disjoint: same axis, other index
            if a in fixed and fixed[a] != i:
                return
        e_free = [a for a in self.names if a not in e_fixed]
        sel = []
        for k, a in enumerate(e_free):    # entry axes that the region fixes
            if a in fixed:
                sel.append(0 if e_values.shape[k] == 1 else fixed[a])
            else:
                sel.append(slice(None))
        sub = e_values[tuple(sel)]
        tsel = tuple(e_fixed[a] if a in e_fixed else slice(None) for a in free)
        target = out[tsel]
        if replace:
            out[tsel] = sub
        elif target.ndim == 0:
            if not np.isnan(sub):
                out[tsel] = sub
        else:
            _merge_into(target, sub)

    def _compose(self, fixed, out):
        free = self.free_axes(fixed)
        for e in self.log:
            self._replay(e.fixed, e.values, False, fixed, free, out)
        for t, level in self.levels.items():
            self._replay({self.time_axis: t}, level, True, fixed, free, out)
        return out

    # -- read ----------------------------------------------------------------
    def read(self, location):
        """Materialised copy of the region ``location`` (free axes in model order)."""
        fixed = self.normalise(location)
        if self.dense is not None:
            return np.array(self.dense[self._index(fixed)], dtype=np.float64)
        t = fixed.get(self.time_axis)
        if t is not None and t in self.levels:
            rest = {a: i for a, i in fixed.items() if a != self.time_axis}
            spatial = [a for a in self.names if a != self.time_axis]
            return np.array(self.levels[t][tuple(rest[a] if a in rest else slice(None) for a in spatial)])
        return self._compose(fixed, np.full(self.region_shape(fixed), np.nan))

    def direct_level(self, t):
        """The stored array of time level ``t`` when no composition is needed
        (a single complete write covers it), else None.  Avoids a host copy of
        large initial conditions."""
        if self.dense is not None:
            return self.dense[self._index({self.time_axis: int(t)})]
        if t in self.levels:
            return self.levels[t]
        spatial = self.region_shape({self.time_axis: t})
        hits = [e for e in self.log
                if self.time_axis not in e.fixed or e.fixed[self.time_axis] == t]
        if len(hits) == 1 and hits[0].fixed == {self.time_axis: t} and hits[0].values.shape == spatial:
            return hits[0].values
        return None

    def uniform_level(self, t):
        """The single value of time level ``t`` when one broadcast write (an
        expression without axis names, e.g. set_IC("0")) is all it holds."""
        if self.dense is not None or t in self.levels:
            return None
        hits = [e for e in self.log
                if self.time_axis not in e.fixed or e.fixed[self.time_axis] == t]
        if len(hits) == 1 and hits[0].fixed == {self.time_axis: t} and hits[0].values.size == 1:
            return float(hits[0].values.reshape(()))
        return None

    # -- write ---------------------------------------------------------------
    def write(self, location, values, copy=True):
        """Merge ``values`` (NaN = leave untouched) into the region."""
        fixed = self.normalise(location)
        values = np.asarray(values) if not copy else np.asarray(values, dtype=np.float64)
        if self.dense is not None:
            idx = self._index(fixed)
            target = self.dense[idx]
            if target.ndim == 0:
                if not np.isnan(values.reshape(())):
                    self.dense[idx] = values.reshape(())
            else:
                _merge_into(target, values)
            return
        nd = len(self.region_shape(fixed))
        if values.ndim != nd:
            values = values.reshape((1,) * (nd - values.ndim) + values.shape)
        self.log.append(Entry(fixed, np.array(values) if copy else values))

    def write_level(self, t, array):
        """Store a complete time level (a snapshot coming back from the device):
        it replaces, NaNs included, what was there."""
        if self.dense is not None:
            self.dense[self._index({self.time_axis: int(t)})] = array
        else:
            self.levels[int(t)] = array

    # -- dense form ------------------------------------------------------------
    def materialise(self):
        """Switch to the dense form (the reference's ``Field.data`` array)."""
        if self.dense is None:
            self.dense = self._compose({}, np.full(self.shape, np.nan))
            self.log = []
            self.levels = {}
        return self.dense

    def time_dependent(self, location, t_from):
        """Does the region (a boundary face) change over time levels >= t_from?"""
        fixed = self.normalise(location)
        if self.dense is not None:
            face = self.dense[self._index(fixed)]
            face = np.moveaxis(face, self.free_axes(fixed).index(self.time_axis), 0)[t_from:]
            return any(not np.array_equal(face[0], f, equal_nan=True) for f in face[1:])
        if any(t >= t_from for t in self.levels):
            return True
        for e in self.log:
            if any(a in fixed and fixed[a] != i for a, i in e.fixed.items()):
                continue
            if self.time_axis in e.fixed:
                if e.fixed[self.time_axis] >= t_from:
                    return True
                continue
            e_free = [a for a in self.names if a not in e.fixed]
            if e.values.shape[e_free.index(self.time_axis)] != 1:
                return True
        return False
