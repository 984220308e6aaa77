"""nfcore -- Python binding of libnfcore (B200 planar-flow free-energy kernels)."""
from ._lib import NFError, load, declared_symbols, LIB_PATH  # noqa: F401
from . import semantics  # noqa: F401
