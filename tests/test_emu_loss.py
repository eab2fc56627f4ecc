"""Kernel-logic check of csrc/loss.cu on the GPU-less build box (host-fiber emulator, tests/emu)."""
import pytest

import emu_backend


@pytest.fixture(scope="module", autouse=True)
def emulated_kernels():
    from msgat_b200 import _lib
    saved = (_lib._LIB, _lib._DEBUG_BACKEND)
    emu_backend.install()
    yield
    _lib._LIB, _lib._DEBUG_BACKEND = saved


def test_emulated_loss_and_metrics_match_oracle():
    import torch
    from loss_cases import run_loss_case
    errs = run_loss_case(torch.device("cpu"))
    bad = {k: v for k, v in errs.items() if not v < 2e-6}
    assert not bad, errs


def test_emulated_adam_matches_torch_adam():
    import torch
    from loss_cases import run_adam_case
    assert run_adam_case(torch.device("cpu")) < 1e-5


def test_emulated_static_scratch_loss_matches_and_rejects_oversize():
    """msgat_loss_forward_static (no counterpart in the reference): same numbers as msgat_loss_forward on two calls
    in a row, and a clear error beyond its size limit.  The GPU check proper is tests/test_zz_no_alloc.py."""
    import ctypes as C
    import torch
    from msgat_b200 import _lib
    lib = _lib.get()
    g = torch.Generator().manual_seed(5)
    for _ in range(2):
        y = torch.rand(3, 37, 12, generator=g) * 300.0
        p = y + torch.randn(3, 37, 12, generator=g) * 60.0
        outs = []
        for static in (False, True):
            loss, d, m = torch.zeros(1), torch.zeros_like(p), torch.zeros(4, dtype=torch.float64)
            if static:
                rc = lib.msgat_loss_forward_static(_lib.ptr(p), _lib.ptr(y), p.numel(), 50.0, 0.0, _lib.ptr(loss), _lib.ptr(d),
                                                   _lib.ptr(m), None)
            else:
                sc = torch.zeros(int(lib.msgat_loss_scratch_bytes()) // 8 + 1, dtype=torch.float64)
                rc = lib.msgat_loss_forward(_lib.ptr(p), _lib.ptr(y), p.numel(), 50.0, 0.0, _lib.ptr(loss), _lib.ptr(d),
                                            _lib.ptr(m), _lib.ptr(sc), None)
            assert rc == 0
            outs.append((loss, d, m))
        for a, b in zip(*outs):
            assert torch.equal(a, b)
    x = torch.zeros(4)
    rc = lib.msgat_loss_forward_static(_lib.ptr(x), _lib.ptr(x), C.c_int64(int(lib.msgat_loss_static_max_n()) + 1), 1.0, 0.0,
                                       _lib.ptr(x), None, None, None)
    assert rc == -63 and b"limit" in lib.msgat_last_error()
