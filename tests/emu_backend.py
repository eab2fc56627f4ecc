"""Inject the host-fiber debug build of the kernels (tests/emu) as the library behind
msgat_b200._lib, so that the kernel logic can be checked against the oracle on the
GPU-less build box.  Tests only."""
import ctypes
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "emu"))
sys.path.insert(0, os.path.dirname(HERE))

import build_emu  # noqa: E402

from msgat_b200 import _lib  # noqa: E402


def install():
    path = build_emu.build()
    lib = ctypes.CDLL(path)
    _lib._declare(lib)
    _lib._LIB = lib
    _lib._DEBUG_BACKEND = True
    return lib
