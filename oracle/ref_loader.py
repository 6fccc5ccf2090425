"""Load the UNMODIFIED reference ``FD.py`` from ``/root/reference`` (build
container only; the GPU box does not have it).

TEST INFRASTRUCTURE ONLY.  Used by ``oracle/make_golden.py`` to generate the
fixtures under ``tests/golden/`` and by ``tests/test_oracle.py`` to validate
the numpy restatement against the live reference when it is present.

The reference hard-imports matplotlib, ipywidgets and numexpr (FD.py:2,4,9),
none of which is installed here.  Empty stand-in modules are registered for
the two plotting libraries (only the out-of-scope plotting methods touch
them) and a ``numexpr`` stand-in whose ``evaluate`` is a restricted ``eval``
over the eleven numpy functions the reference whitelists (FD.py:405).
"""

from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np

REFERENCE_FD = "/root/reference/src/finite_difference/FD.py"

_MATH = ("exp", "abs", "sin", "cos", "tan", "arcsin", "arccos", "arctan",
         "sinh", "cosh", "tanh")


def available() -> bool:
    return os.path.isfile(REFERENCE_FD)


def _numexpr_stub():
    mod = types.ModuleType("numexpr")

    def evaluate(expr, local_dict=None, **_):
        env = {n: getattr(np, n) for n in _MATH}
        env["__builtins__"] = {}
        loc = {k: np.asarray(v) for k, v in (local_dict or {}).items()}
        return np.asarray(eval(expr, env, loc))

    mod.evaluate = evaluate
    return mod


def load():
    """Return the reference module (cached)."""
    if "_fd_reference" in sys.modules:
        return sys.modules["_fd_reference"]
    if not available():
        raise FileNotFoundError(REFERENCE_FD)
    saved = {}
    stubs = {
        "matplotlib": types.ModuleType("matplotlib"),
        "matplotlib.pyplot": types.ModuleType("matplotlib.pyplot"),
        "ipywidgets": types.ModuleType("ipywidgets"),
        "ipywidgets.widgets": types.ModuleType("ipywidgets.widgets"),
        "numexpr": _numexpr_stub(),
    }
    stubs["matplotlib"].pyplot = stubs["matplotlib.pyplot"]
    stubs["ipywidgets"].widgets = stubs["ipywidgets.widgets"]
    for name, mod in stubs.items():
        saved[name] = sys.modules.get(name)
        sys.modules[name] = mod
    try:
        spec = importlib.util.spec_from_file_location("_fd_reference", REFERENCE_FD)
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
    finally:
        for name, old in saved.items():
            if old is None:
                sys.modules.pop(name, None)
            else:
                sys.modules[name] = old
    sys.modules["_fd_reference"] = ref
    return ref
