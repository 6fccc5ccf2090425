"""``Model.interact`` (FD.py:118-216): notebook slider over ``Field.data``.

Out of scope as compute: it only reads ``Field.data`` and ``Model.dt``.  The
plotting libraries are imported lazily; the slicing helper is plain numpy so
it can be tested without them.
"""

import numpy as np


def nearest_time_slice(model, data, t):
    """Slice of ``data`` at the time level closest to ``t`` (FD.py:178-186)."""
    values = model.axes[model.time_axis].tolist()
    diffs = [abs(t - v) for v in values]
    sel = [slice(None)] * data.ndim
    sel[model.axes_names.index(model.time_axis)] = diffs.index(min(diffs))
    return data[tuple(sel)]


def value_range(data):
    """Colour / y range over all time levels (FD.py:168-176)."""
    hi, lo = np.nanmax(data), np.nanmin(data)
    return (0 if np.isnan(lo) else lo), (0.1 if np.isnan(hi) else hi)


def interact(model):
    assert model.time_axis is not None, "need time axis to use interact"
    assert len(model.fields) != 0, "model has no fields"
    try:
        import ipywidgets.widgets as widgets
        import matplotlib.pyplot as plt
    except ImportError as exc:
        raise ImportError("interact needs matplotlib and ipywidgets") from exc

    fields = list(reversed(model.fields))   # derivatives after their fields
    names = [f.name for f in fields]
    space = {a: v for a, v in model.axes.items() if a != model.time_axis}
    if len(space) > 2:
        raise Exception("for now can only visualize \
                            with 1 or 2 space dimensions")
    times = model.axes[model.time_axis]
    plt.close("all")
    _, ax = plt.subplots(figsize=(6, 4))
    space_names, space_values = list(space.keys()), list(space.values())

    @widgets.interact(field_name=widgets.Dropdown(options=names, description='Field'),
                      t=(times[0], times[-1], model.dt))
    def update(field_name=names[0], t=0):
        ax.clear()
        data = fields[names.index(field_name)].data
        lo, hi = value_range(data)
        level = nearest_time_slice(model, data, t)
        x = space_values[0]
        ax.set_xlim(x[0], x[-1])
        ax.set_xlabel(space_names[0])
        if len(space) == 1:
            ax.set_ylabel(field_name)
            ax.set_ylim(lo, hi)
            if len(level) > len(x):     # fixed edge axes carry one extra point
                level = level[:-1]
            ax.plot(x, level, color="blue")
        else:
            y = space_values[1]
            ax.set_ylabel(space_names[1])
            ax.set_ylim(y[0], y[-1])
            ax.imshow(np.flipud(np.transpose(level)), extent=[x[0], x[-1], y[0], y[-1]], vmin=lo, vmax=hi)
