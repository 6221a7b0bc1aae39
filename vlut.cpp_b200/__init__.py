"""vlut.cpp_b200 -- B200-native (sm_100a) Vec-LUT ternary MUL_MAT: CUDA kernels behind a C ABI, plus the host
mirror of the reference's operator interface.  Import name: ``vlut_b200`` (see vlut_b200.py at the repo root; the
directory name carries a dot and cannot be imported directly)."""
from . import _lib  # noqa: F401
from ._lib import LIB_PATH, VlutError, declared_symbols  # noqa: F401


def __getattr__(name):  # lazy: the compute API needs torch; symbol/loader checks do not
    if name in ("api",):
        import importlib
        return importlib.import_module(__name__ + ".api")
    raise AttributeError(name)
