"""Hand-written sm_100a kernels and their autograd wrappers.

The kernels live in ``pipe_b200/csrc`` behind the C ABI of ``include/pipe_b200.h``
and are bound with ctypes.  On a machine with a GPU the native library is
mandatory: ops raise if it is missing instead of silently running eager code.
"""
from __future__ import annotations

from ._lib import NativeUnavailable, available, load


def require_native() -> None:
    """Fail loudly unless the native library loads."""
    load()


def launch_count() -> int:
    """Number of kernels this library has launched in this process."""
    return int(load().pb2_launch_count())
