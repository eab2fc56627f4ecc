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
