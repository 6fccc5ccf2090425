"""The C-ABI shared library: it loads, and exports every symbol that
include/fd_b200.h declares (no compute calls: this runs without a GPU)."""

import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "fd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fd_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import finite_difference._native as nat
    names = declared_symbols()
    assert len(names) >= 17
    lib = ctypes.CDLL(nat.LIB_PATH)
    for name in names:
        assert hasattr(lib, name), f"{name} declared in fd_b200.h but not exported"
        assert name in nat.SIGNATURES, f"{name} has no ctypes signature in _native.py"
    assert lib.fd_abi_version() == nat.ABI_VERSION
    assert set(nat.SIGNATURES) == set(names)


def test_struct_sizes_match_the_header():
    """ctypes mirrors of the descriptor structs must have the C layout."""
    import subprocess
    import tempfile
    import finite_difference._native as nat
    src = '#include <stdio.h>\n#include "fd_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(fd_stencil_t), sizeof(fd_term_t), sizeof(fd_post_t), sizeof(fd_fused_t), sizeof(fd_pair_axis_t), sizeof(fd_pair_t), sizeof(fd_expr_t));return 0;}\n'
    with tempfile.TemporaryDirectory() as tmp:
        c = os.path.join(tmp, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(tmp, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        sizes = [int(v) for v in subprocess.check_output([exe]).split()]
    assert sizes == [ctypes.sizeof(nat.Stencil), ctypes.sizeof(nat.Term), ctypes.sizeof(nat.Post),
                     ctypes.sizeof(nat.Fused), ctypes.sizeof(nat.PairAxis), ctypes.sizeof(nat.Pair), ctypes.sizeof(nat.Expr)]


def test_missing_device_fails_loudly():
    """No CPU fallback: advancing a model without a CUDA device raises."""
    import contextlib
    import io
    import numpy as np
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import finite_difference as fd
    from finite_difference._device import NoDeviceError
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model({"t": range(3), "x": range(8)}, "t", periodic=("x",))
        T = fd.Field(m, "T", n_time_ders=1)
        T.set_IC("x")
        d = fd.Stencil([-1, 0, 1], 2)
        T.dot.assign_update(d.der(T.prev, "x"))
        T.time_integrate_update()
    with pytest.raises(NoDeviceError):
        m.increment_time()
