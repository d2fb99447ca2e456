"""b200_sparse: ctypes host layer over libb200sparse.so (sm_100a kernels, C ABI)."""
from . import cabi, ops  # noqa: F401
from .cabi import ABI_VERSION, B200Error, LIB_PATH  # noqa: F401
