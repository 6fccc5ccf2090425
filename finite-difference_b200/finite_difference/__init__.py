"""finite_difference -- B200-native implementation of the time-stepping hot
path of Ephraim-Bryski/Finite-Difference 0.0.3, behind the unchanged public
API (``Model``, ``Stencil``, ``Field``, ``ConstantField``; reference
``__init__.py:1-3``).  Host code is Python; every array operation of a time
step runs in hand-written sm_100a CUDA kernels (``libfd_b200.so``) reached
through ctypes.  There is no CPU fallback.
"""

from . import _native  # noqa: F401  (fails loudly if libfd_b200.so is missing)
from ._graph import FieldInstant
from .field import ConstantField, Field
from .model import Model
from .stencil import Stencil

__all__ = ["Model", "Stencil", "Field", "ConstantField"]
__version__ = "0.0.3+b200.1"
