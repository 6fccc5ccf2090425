"""IC/BC expression strings -> arrays of axis coordinates (host side, set-up only).

Replaces the reference's ``itertools.product`` + ``numexpr.evaluate`` pair
(FD.py:446-471), which builds O(points) Python objects, by a broadcast numpy
evaluation with the same grammar, whitelist and error messages.  The result is
returned *un-broadcast*: axes the expression does not mention keep extent 1,
so that a boundary face "0" for 1000 time levels costs 8 bytes.
"""

from __future__ import annotations

import re

import numpy as np

# FD.py:405 -- the only non-axis identifiers an expression may use
MATH_TERMS = ["exp", "abs", "sin", "cos", "tan", "arcsin", "arccos", "arctan", "sinh", "cosh", "tanh"]
_IDENT = re.compile(r"\b[A-Za-z_][A-Za-z0-9_]*\b")  # FD.py:404
_ENV = {name: getattr(np, name) for name in MATH_TERMS}
_ENV["abs"] = np.abs
_ENV["__builtins__"] = {}


def variables_of(expression):
    return [v for v in _IDENT.findall(expression) if v not in MATH_TERMS]


def evaluate(expression, all_axes, free_axes, coords):
    """Evaluate ``expression`` over the tensor product of ``coords``.

    all_axes: every axis name of the model; free_axes: the axes the region
    spans, in model order; coords: their sample coordinates.  Returns a
    float64 array with ``len(free_axes)`` dimensions whose extent is 1 along
    axes the expression does not depend on.
    """
    variables = variables_of(expression)
    unknown_vars = set(variables) - set(all_axes)
    if len(unknown_vars) != 0:
        raise Exception(f"variables {unknown_vars} are not axes names")  # FD.py:451
    unsubbed_vars = set(variables) - set(free_axes)
    if len(unsubbed_vars) != 0:
        raise Exception(  # FD.py:454
            f"variables {unsubbed_vars} are axes names along which the expression is being set")
    nd = len(free_axes)
    local = {}
    for k, (name, c) in enumerate(zip(free_axes, coords)):
        if name in variables:
            shape = [1] * nd
            shape[k] = len(c)
            local[name] = np.asarray(c).reshape(shape)
    try:
        with np.errstate(all="ignore"):
            value = eval(compile(expression, "<expression>", "eval"), dict(_ENV), local)
        value = np.asarray(value, dtype=np.float64)
    except Exception:
        raise Exception("cannot parse expression")  # FD.py:464
    if value.ndim == 0:
        value = value.reshape([1] * nd)
    if value.ndim != nd:
        raise Exception("cannot parse expression")
    return value
