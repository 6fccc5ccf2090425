"""Load the UNMODIFIED reference ``FD.py``: from ``/root/reference`` in the
build container, or from ``oracle/_ref/FD.py`` where ``__graft_entry__.build()``
staged a copy (git-ignored, never committed; it travels to the GPU box with
the working tree like the built ``.so``).

TEST INFRASTRUCTURE ONLY.  Used by ``oracle/make_golden.py`` to generate the
fixtures under ``tests/golden/``, by ``tests/test_oracle.py`` to validate the
numpy restatement against the live reference when it is present, and by the
CPU legs of ``bench.py`` (``cpu_baseline``, ``--impl reference``), which time it.

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
STAGED_FD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "FD.py")

_MATH = ("exp", "abs", "sin", "cos", "tan", "arcsin", "arccos", "arctan",
         "sinh", "cosh", "tanh")


def location():
    """Path of the reference source to load, or None."""
    for path in (REFERENCE_FD, STAGED_FD):
        if os.path.isfile(path):
            return path
    return None


def available() -> bool:
    return location() is not None


def stage() -> bool:
    """Copy the reference file next to the oracle (build container only)."""
    if not os.path.isfile(REFERENCE_FD):
        return os.path.isfile(STAGED_FD)
    import shutil
    os.makedirs(os.path.dirname(STAGED_FD), exist_ok=True)
    shutil.copyfile(REFERENCE_FD, STAGED_FD)
    return True


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
    path = location()
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
        spec = importlib.util.spec_from_file_location("_fd_reference", path)
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
