"""ctypes binding of ``libfd_b200.so`` (declared in ``include/fd_b200.h``).

There is no CPU fallback: if the shared library is missing the import of this
module fails loudly and nothing in the package can execute a time step.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# (FD_B200_LIB: another build of the same library, for A/B runs of kernel variants)
LIB_PATH = os.environ.get("FD_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libfd_b200.so")

ABI_VERSION = 6
FD_MAX_POINTS = 16
FD_MAX_TERMS = 4
FD_MAX_POST = 4
FD_FLAG_INF = 0
FD_FLAG_OVERRIDE = 1
FD_FLAG_HALO_TIMEOUT = 2
FD_IPC_HANDLE_BYTES = 64
FD_NUM_FLAGS = 4

# enum fd_ewise_op
(OP_NEG, OP_ADD_FF, OP_SUB_FF, OP_MUL_FF, OP_DIV_FF, OP_POW_FF, OP_ADD_FS, OP_SUB_FS,
 OP_MUL_FS, OP_DIV_FS, OP_POW_FS, OP_SQUARE, OP_RECIP, OP_SQRT, OP_COPY, OP_EULER) = range(16)

# enum fd_post_op
POST_MUL, POST_ADD, POST_SUB, POST_DIV, POST_NEG, POST_MUL_FIELD = range(6)


class Stencil(C.Structure):
    _fields_ = [("npts", C.c_int32),
                ("shifts", C.c_int32 * FD_MAX_POINTS),
                ("weights", C.c_double * FD_MAX_POINTS),
                ("scale", C.c_double),
                ("exact_div", C.c_int32)]


class Term(C.Structure):
    _fields_ = [("src", C.c_void_p),
                ("axis", C.c_int32), ("periodic", C.c_int32),
                ("lo", C.c_int32), ("hi", C.c_int32),
                ("n_src", C.c_int32), ("subtract", C.c_int32),
                ("has_pre", C.c_int32), ("negate", C.c_int32),
                ("fold_div", C.c_int32), ("fold_mod", C.c_int32),
                ("pre", C.c_double),
                ("st", Stencil)]


class Post(C.Structure):
    _fields_ = [("op", C.c_int32), ("_pad", C.c_int32), ("s", C.c_double), ("coef", C.c_void_p)]


class Fused(C.Structure):
    _fields_ = [("n", C.c_int32 * 3),
                ("nterms", C.c_int32),
                ("term", Term * FD_MAX_TERMS),
                ("npost", C.c_int32), ("depth", C.c_int32),
                ("post", Post * FD_MAX_POST),
                ("dt", C.c_double),
                ("dot_out", C.c_void_p),
                ("prev", C.c_void_p * 2),
                ("out", C.c_void_p * 2),
                ("face", ((C.c_void_p * 2) * 3) * 2),
                ("nan_count", C.c_void_p * 2),
                ("flags", C.c_void_p),
                ("box_lo", C.c_int32 * 3), ("box_hi", C.c_int32 * 3),
                ("rim_lo", C.c_int32 * 3), ("rim_hi", C.c_int32 * 3)]


class PairAxis(C.Structure):
    _fields_ = [("v_prev", C.c_void_p), ("v_out", C.c_void_p), ("vdot_out", C.c_void_p),
                ("v_nan_count", C.c_void_p), ("c_up", C.c_double),
                ("axis", C.c_int32), ("periodic", C.c_int32), ("subtract", C.c_int32), ("_pad", C.c_int32),
                ("up", Stencil), ("down", Stencil)]


class Pair(C.Structure):
    _fields_ = [("n", C.c_int32 * 3), ("naxes", C.c_int32),
                ("ax", PairAxis * 3),
                ("p_prev", C.c_void_p), ("p_out", C.c_void_p), ("pdot_out", C.c_void_p),
                ("p_nan_count", C.c_void_p), ("flags", C.c_void_p),
                ("c_down", C.c_double), ("dt", C.c_double),
                ("box_lo", C.c_int32 * 3), ("box_hi", C.c_int32 * 3)]


FD_EXPR_MAX_FIELDS, FD_EXPR_MAX_CODE, FD_EXPR_MAX_STACK = 4, 32, 8
# enum fd_expr_op
(X_PUSH_DER, X_PUSH_FIELD, X_PUSH_IMM, X_NEG, X_ADD, X_SUB, X_MUL, X_DIV, X_POW, X_ADD_IMM, X_SUB_IMM, X_MUL_IMM,
 X_DIV_IMM, X_POW_IMM, X_SQUARE, X_RECIP, X_SQRT) = range(17)


class Expr(C.Structure):
    _fields_ = [("base", Fused),
                ("nfields", C.c_int32), ("ncode", C.c_int32),
                ("field", C.c_void_p * FD_EXPR_MAX_FIELDS),
                ("op", C.c_int32 * FD_EXPR_MAX_CODE),
                ("arg", C.c_int32 * FD_EXPR_MAX_CODE),
                ("imm", C.c_double * FD_EXPR_MAX_CODE)]


FD_MAX_LINKS = 4


class HaloLink(C.Structure):
    _fields_ = [("src_lo", C.c_void_p), ("peer_dst_lo", C.c_void_p), ("src_hi", C.c_void_p), ("peer_dst_hi", C.c_void_p),
                ("nbytes", C.c_longlong), ("peer_flag_lo", C.c_void_p), ("peer_flag_hi", C.c_void_p),
                ("my_flag_lo", C.c_void_p), ("my_flag_hi", C.c_void_p), ("value", C.c_longlong)]


class SlabStep(C.Structure):
    _fields_ = [("kind", C.c_int32), ("dim", C.c_int32), ("lo", C.c_int32), ("hi", C.c_int32),
                ("depth", C.c_int32), ("nlinks", C.c_int32), ("desc", C.c_void_p),
                ("link", HaloLink * FD_MAX_LINKS), ("timeout_ns", C.c_longlong), ("flags", C.c_void_p),
                ("compute_stream", C.c_void_p), ("comm_stream", C.c_void_p)]


_I64, _P, _D, _I = C.c_int64, C.c_void_p, C.c_double, C.c_int

# name -> (restype, argtypes); every symbol include/fd_b200.h declares
SIGNATURES = {
    "fd_abi_version": (C.c_int, []),
    "fd_last_error": (C.c_char_p, []),
    "fd_launch_count": (C.c_longlong, []),
    "fd_last_kernel": (C.c_char_p, []),
    "fd_ipc_export": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_longlong)]),
    "fd_ipc_open": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "fd_ipc_close": (C.c_int, [C.c_void_p]),
    "fd_halo_push": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p,
                               C.c_longlong, C.c_void_p]),
    "fd_halo_wait": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_longlong, C.c_void_p, C.c_void_p]),
    "fd_halo_exchange": (C.c_int, [C.POINTER(HaloLink), C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]),
    "fd_seq_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "fd_seq_destroy": (C.c_int, [C.c_void_p]),
    "fd_seq_wait_halo": (C.c_int, [C.c_void_p, C.c_void_p]),
    "fd_slab_step_f64": (C.c_int, [C.c_void_p, C.POINTER(SlabStep)]),
    "fd_slab_step_f32": (C.c_int, [C.c_void_p, C.POINTER(SlabStep)]),
    "fd_fused_step_boxes_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fd_fused_step_boxes_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
}
SIGNATURES["fd_fused_nested_static_f64"] = (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p])
FD_STATIC_SCRATCH_BYTES = 16 * 1024 * 1024
for _sfx in ("f64", "f32"):
    SIGNATURES.update({
        f"fd_der_{_sfx}": (_I, [_P, _P, _I64, _I64, _I64, _I64, C.POINTER(Stencil), _I, _I64, _I64, _P, _P]),
        f"fd_ewise_{_sfx}": (_I, [_I, _P, _P, _P, _D, _I64, _P]),
        f"fd_merge_store_{_sfx}": (_I, [_P, _P, _I64, _P, _P, _P]),
        f"fd_fill_{_sfx}": (_I, [_P, _D, _I64, _P]),
        f"fd_face_store_{_sfx}": (_I, [_P, _P, _I64, _I64, _I64, _I64, _P]),
        f"fd_count_nan_{_sfx}": (_I, [_P, _I64, _P, _P]),
        f"fd_fused_step_{_sfx}": (_I, [C.POINTER(Fused), _P]),
        f"fd_fused_pair_{_sfx}": (_I, [C.POINTER(Pair), _P]),
        f"fd_fused_expr_{_sfx}": (_I, [C.POINTER(Expr), _P]),
        f"fd_fused_nested_{_sfx}": (_I, [_I, _P, _I, _P, _P, C.c_longlong, _P]),
    })


class NativeError(RuntimeError):
    pass


def _load():
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"finite_difference (B200): {LIB_PATH} is missing. Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` from the repository root "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.fd_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {lib.fd_abi_version()}, expected {ABI_VERSION}")
    return lib


lib = _load()


def check(rc):
    if rc != 0:
        raise NativeError(lib.fd_last_error().decode())


def suffix(dtype) -> str:
    import numpy as np
    return "f32" if np.dtype(dtype) == np.float32 else "f64"


def fn(name, dtype):
    return getattr(lib, f"{name}_{suffix(dtype)}")


def make_stencil(shifts, weights, scale, exact_div=True) -> Stencil:
    if len(shifts) > FD_MAX_POINTS:
        raise NativeError(f"stencils with more than {FD_MAX_POINTS} points are not supported by the B200 kernels")
    st = Stencil()
    st.npts = len(shifts)
    for i, (s, w) in enumerate(zip(shifts, weights)):
        st.shifts[i] = int(s)
        st.weights[i] = float(w)
    st.scale = float(scale)
    st.exact_div = 1 if exact_div else 0
    return st
