"""Kernel layer: ctypes bindings of libppb200.so (sm_100a) + fp32 torch references."""
from . import _lib, functional, reference  # noqa: F401
