"""Shared constants and the error type (cf. python/dgl/base.py and _ffi/base.py:DGLError)."""
import warnings

ALL = "__ALL__"


class DGLError(Exception):
    """Error raised by the engine; same name as the reference's dgl._ffi.base.DGLError."""


def is_all(arg):
    return isinstance(arg, str) and arg == ALL


def dgl_warning(msg):
    warnings.warn(msg)
