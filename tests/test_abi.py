"""The C-ABI library loads and exports every symbol include/msgat_sm100.h declares
(no compute: the build box has no GPU)."""
import os
import re

from conftest import ROOT

from msgat_b200 import _lib


def declared_symbols():
    with open(os.path.join(ROOT, "include", "msgat_sm100.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(msgat_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert sorted(_lib.EXPORTS) == declared_symbols()


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    lib = _lib.load_unchecked()
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.msgat_version()


def test_no_fallback_without_device():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    saved, _lib._LIB = _lib._LIB, None
    try:
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            _lib.get()
    finally:
        _lib._LIB = saved


def test_struct_sizes_match_header():
    import ctypes
    assert ctypes.sizeof(_lib.MeamDesc) == 4 * 8 + 4 * 4 + 4 * 2 + 8 * 3
    assert ctypes.sizeof(_lib.MeamParams) == 8 * (11 + 8 + 3)
    assert ctypes.sizeof(_lib.Adj) == 8 + 8 * 6
    assert ctypes.sizeof(_lib.HeadDesc) == 4 * 6 + 8
