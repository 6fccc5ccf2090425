"""API misuse cases: each snippet must raise the same
exception type with the same message in the B200 package as in the reference.

TEST INFRASTRUCTURE ONLY.  ``CASES`` maps a name to a function ``f(fd)`` that
is expected to raise; ``oracle/make_golden.py`` records ``(type, message)`` of
the reference for each into ``tests/golden/misuse.json``.
"""

import numpy as np


def _model(fd, **kw):
    return fd.Model({"t": range(0, 4), "x": range(0, 6)}, "t", **kw)


def _setup(fd):
    m = _model(fd)
    T = fd.Field(m, "T", n_time_ders=1)
    d2 = fd.Stencil([-1, 0, 1], 2)
    return m, T, d2


CASES = {}


def case(fn):
    CASES[fn.__name__] = fn
    return fn


# ---- Model -----------------------------------------------------------------
@case
def axes_not_dict(fd):
    fd.Model([("t", [0, 1])], "t")


@case
def time_axis_not_str(fd):
    fd.Model({"t": [0, 1]}, 3)


@case
def periodic_wrong_type(fd):
    fd.Model({"t": [0, 1], "x": [0, 1]}, "t", periodic="x")


@case
def axis_key_not_str(fd):
    fd.Model({"t": [0, 1], 1: [0, 1]}, "t")


@case
def axis_not_iterable(fd):
    fd.Model({"t": [0, 1], "x": 5}, "t")


@case
def axis_not_numbers(fd):
    fd.Model({"t": [0, 1], "x": ["a", "b"]}, "t")


@case
def axis_too_short(fd):
    fd.Model({"t": [0, 1], "x": [1]}, "t")


@case
def axis_uneven(fd):
    fd.Model({"t": [0, 1], "x": [0, 1, 3]}, "t")


@case
def axis_descending(fd):
    fd.Model({"t": [0, 1], "x": [3, 2, 1]}, "t")


@case
def time_axis_missing(fd):
    fd.Model({"t": [0, 1]}, "s")


@case
def time_axis_periodic(fd):
    fd.Model({"t": [0, 1], "x": [0, 1]}, "t", periodic=("t",))


@case
def periodic_unknown_axis(fd):
    fd.Model({"t": [0, 1], "x": [0, 1]}, "t", periodic=("q",))


@case
def check_ic_missing(fd):
    m, T, d2 = _setup(fd)
    m.check_IC()


@case
def increment_without_update(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    m.increment_time()


# ---- Field -----------------------------------------------------------------
@case
def field_bad_model(fd):
    fd.Field("m", "T")


@case
def field_bad_name(fd):
    fd.Field(_model(fd), 3)


@case
def field_bad_edge_type(fd):
    fd.Field(_model(fd), "T", edge_axes=3)


@case
def field_unknown_edge_axis(fd):
    fd.Field(_model(fd), "T", edge_axes=("q",))


@case
def field_time_edge_axis(fd):
    fd.Field(_model(fd), "T", edge_axes=("t",))


@case
def field_bad_ders(fd):
    fd.Field(_model(fd), "T", n_time_ders=-1)


@case
def new_before_update(fd):
    m, T, d2 = _setup(fd)
    T.new


@case
def prev_after_update(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.dot.assign_update(d2.der(T.prev, "x"))
    T.time_integrate_update()
    T.prev


@case
def prev_without_dot(fd):
    m = _model(fd)
    q = fd.Field(m, "q")
    q.prev


@case
def set_ic_without_dot(fd):
    m = _model(fd)
    fd.Field(m, "q").set_IC("1")


@case
def set_bc_time_axis(fd):
    m, T, d2 = _setup(fd)
    T.set_BC("0", "t", "start")


@case
def set_bc_unknown_axis(fd):
    m, T, d2 = _setup(fd)
    T.set_BC("0", "q", "start")


@case
def set_bc_periodic_axis(fd):
    m = _model(fd, periodic=("x",))
    fd.Field(m, "T", n_time_ders=1).set_BC("0", "x", "start")


@case
def set_bc_bad_side(fd):
    m, T, d2 = _setup(fd)
    T.set_BC("0", "x", "left")


@case
def expression_not_str(fd):
    m, T, d2 = _setup(fd)
    T.set_IC(1)


@case
def expression_unknown_variable(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("sqrt(x)")


@case
def expression_uses_fixed_axis(fd):
    m, T, d2 = _setup(fd)
    T.set_BC("x", "x", "start")


@case
def expression_garbage(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x +* 2")


@case
def assign_twice(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.dot.assign_update(d2.der(T.prev, "x"))
    T.dot.assign_update(d2.der(T.prev, "x"))


@case
def assign_on_integrated_field(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.assign_update(T.prev)


@case
def assign_not_an_instant(fd):
    m, T, d2 = _setup(fd)
    T.dot.assign_update(np.zeros(6))


@case
def assign_edge_mismatch(fd):
    m = _model(fd)
    T = fd.Field(m, "T", n_time_ders=1)
    T.set_IC("x")
    ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
    T.dot.assign_update(ce.der(T.prev, "x"))


@case
def integrate_twice(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.dot.assign_update(d2.der(T.prev, "x"))
    T.time_integrate_update()
    T.time_integrate_update()


@case
def integrate_without_dot(fd):
    m = _model(fd)
    fd.Field(m, "q").time_integrate_update()


@case
def integrate_before_dot_update(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.time_integrate_update()


@case
def field_arithmetic(fd):
    m, T, d2 = _setup(fd)
    -T


@case
def field_add(fd):
    m, T, d2 = _setup(fd)
    T + 1


@case
def field_sub_arity(fd):
    m, T, d2 = _setup(fd)
    T - 1


# ---- ConstantField ------------------------------------------------------------
@case
def constant_prev(fd):
    fd.ConstantField(_model(fd), "k").prev


@case
def constant_set_ic(fd):
    fd.ConstantField(_model(fd), "k").set_IC("1")


@case
def constant_set_bc(fd):
    fd.ConstantField(_model(fd), "k").set_BC("1")


# ---- FieldInstant ---------------------------------------------------------------
@case
def instant_plus_field(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.prev + T


@case
def instant_plus_string(fd):
    m, T, d2 = _setup(fd)
    T.set_IC("x")
    T.prev + "a"


@case
def instant_other_model(fd):
    m, T, d2 = _setup(fd)
    m2, T2, _ = _setup(fd)
    T.prev + T2.prev


@case
def instant_edge_mismatch(fd):
    m = _model(fd)
    a = fd.Field(m, "a", n_time_ders=1)
    b = fd.Field(m, "b", edge_axes=("x",), n_time_ders=1)
    a.prev + b.prev


@case
def instant_divide_by_field(fd):
    m, T, d2 = _setup(fd)
    T.prev / T


@case
def instant_pow_field(fd):
    m, T, d2 = _setup(fd)
    T.prev ** T


@case
def instant_rtruediv(fd):
    m, T, d2 = _setup(fd)
    2 / T.prev


@case
def instant_rpow(fd):
    m, T, d2 = _setup(fd)
    2 ** T.prev


@case
def instant_times_array(fd):
    m, T, d2 = _setup(fd)
    T.prev * np.ones(6)


# ---- Stencil ----------------------------------------------------------------------
@case
def stencil_points_not_list(fd):
    fd.Stencil((-1, 0, 1), 2)


@case
def stencil_bad_order(fd):
    fd.Stencil([-1, 0, 1], 0)


@case
def stencil_duplicate_points(fd):
    fd.Stencil([0, 0, 1], 1)


@case
def stencil_bad_axis_type(fd):
    fd.Stencil([-1, 0, 1], 1, "face")


@case
def stencil_bad_der_axis_type(fd):
    fd.Stencil([-1, 0, 1], 1, "cell", "face")


@case
def stencil_too_few_points(fd):
    fd.Stencil([0, 1], 2)


@case
def stencil_mixed_points(fd):
    fd.Stencil([0, 0.5, 1.25], 1)


@case
def stencil_edge_to_cell_integer(fd):
    fd.Stencil([-1, 0, 1], 1, "edge", "cell")


@case
def stencil_edge_to_edge_half(fd):
    fd.Stencil([-.5, .5], 1, "edge", "edge")


@case
def stencil_cell_to_cell_half(fd):
    fd.Stencil([-.5, .5], 1)


@case
def der_on_field(fd):
    m, T, d2 = _setup(fd)
    d2.der(T, "x")


@case
def der_axis_not_str(fd):
    m, T, d2 = _setup(fd)
    d2.der(T.prev, 1)


@case
def der_time_axis(fd):
    m, T, d2 = _setup(fd)
    d2.der(T.prev, "t")


@case
def der_missing_axis_argument(fd):
    m, T, d2 = _setup(fd)
    d2.der(T.prev)


@case
def der_cell_stencil_on_edge_field(fd):
    m = _model(fd)
    b = fd.Field(m, "b", edge_axes=("x",), n_time_ders=1)
    fd.Stencil([-1, 0, 1], 2).der(b.prev, "x")


@case
def der_edge_stencil_on_cell_field(fd):
    m, T, d2 = _setup(fd)
    fd.Stencil([-.5, .5], 1, "edge", "cell").der(T.prev, "x")


@case
def finished_past_the_end(fd):
    m, T, d2 = _setup(fd)
    m.time_step = 4
    m.finished
