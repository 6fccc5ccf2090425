"""Drop-in behaviour of the public API on misuse: every
case of oracle/misuse_cases.py must raise the same exception type with the
same message as the unmodified reference did (tests/golden/misuse.json), and
the printed stencil formulas / weights must be the reference's."""

import contextlib
import io
import json
import os

import pytest

import finite_difference as fd
from oracle import misuse_cases

GOLD = os.path.join(os.path.dirname(__file__), "golden")
REFERENCE = json.load(open(os.path.join(GOLD, "misuse.json")))

# set literals print in hash order; compare those messages up to the set's text
_SET_MESSAGES = {"expression_unknown_variable", "expression_uses_fixed_axis", "periodic_unknown_axis"}


@pytest.mark.parametrize("name", sorted(misuse_cases.CASES))
def test_misuse_raises_like_the_reference(name):
    want_type, want_msg = REFERENCE[name]
    with pytest.raises(BaseException) as info:
        with contextlib.redirect_stdout(io.StringIO()):
            misuse_cases.CASES[name](fd)
    got_type, got_msg = type(info.value).__name__, str(info.value)
    assert got_type == want_type, (name, got_type, got_msg, want_msg)
    if name in _SET_MESSAGES:
        assert got_msg.split("{")[0] == want_msg.split("{")[0] and got_msg.split("}")[-1] == want_msg.split("}")[-1]
    else:
        assert got_msg == want_msg, name


def test_stencil_weights_and_text_like_the_reference():
    table = json.load(open(os.path.join(GOLD, "weights.json")))
    for row in table:
        half = any(p % 1 for p in row["points"])
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            st = fd.Stencil(list(row["points"]), row["order"], "cell", "edge" if half else "cell")
        assert list(st.weights) == row["weights"], row
        assert buf.getvalue() == row["text"], row


def test_str_of_field_and_instant():
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model({"t": range(0, 4), "x": range(0, 6)}, "t")
        T = fd.Field(m, "T", n_time_ders=1)
    assert str(T) == "2-dimensional Field named T, dimension lengths: {'t': 4, 'x': 6}"
    assert str(T.prev) == "2-dimensional field at the current time,\nDimension lengths: {'t': 4, 'x': 6}"
    assert [f.name for f in m.fields] == ["T dot", "T"]


def test_interact_slicing_helpers():
    """The numpy part of Model.interact (FD.py:168-186); the plotting stack
    itself is unavailable on this box."""
    import numpy as np
    from finite_difference._interact import nearest_time_slice, value_range
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model({"x": np.arange(5) * 1.0, "t": np.arange(4) * 0.5}, "t")
    data = np.arange(20.0).reshape(5, 4)
    assert np.array_equal(nearest_time_slice(m, data, 0.8), data[:, 2])
    assert np.array_equal(nearest_time_slice(m, data, -3), data[:, 0])
    assert value_range(np.full((2, 2), np.nan)) == (0, 0.1)
    assert value_range(data) == (0.0, 19.0)
    with pytest.raises(ImportError):
        fd.Field(m, "T", n_time_ders=1).plot({"t": 0})
