"""GPU: the allocation-free loss entry point (msgat_loss_forward_static, csrc/loss.cu) against the existing one.

The static-scratch variant has no counterpart in the reference (it was added
); its contract is "same result as msgat_loss_forward, no allocation".  Called through the C ABI twice in a
row (the scratch array keeps the previous call's contents), on two sizes, and with an n beyond its limit.
This file is named so that pytest collects it last.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-6   # the bar of tests/test_gpu_loss.py


def _call_pair(lib, _lib, pred, target, delta, mask):
    dev = pred.device
    n = pred.numel()
    stream = _lib.stream_of(pred)
    # existing entry point: caller-provided scratch
    nb = int(lib.msgat_loss_scratch_bytes())
    scratch = torch.zeros((nb + 7) // 8, dtype=torch.float64, device=dev)
    loss_a = torch.empty(1, device=dev)
    d_a = torch.empty_like(pred)
    m_a = torch.zeros(4, dtype=torch.float64, device=dev)
    _lib.check(lib.msgat_loss_forward(_lib.ptr(pred), _lib.ptr(target), n, delta, mask, _lib.ptr(loss_a), _lib.ptr(d_a),
                                      _lib.ptr(m_a), _lib.ptr(scratch), stream), "msgat_loss_forward")
    # new entry point: outputs allocated up front, nothing allocated by the call itself
    loss_b = torch.full((1,), float("nan"), device=dev)
    d_b = torch.full_like(pred, float("nan"))
    m_b = torch.zeros(4, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
    _lib.check(lib.msgat_loss_forward_static(_lib.ptr(pred), _lib.ptr(target), n, delta, mask, _lib.ptr(loss_b),
                                             _lib.ptr(d_b), _lib.ptr(m_b), stream), "msgat_loss_forward_static")
    assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before
    torch.cuda.synchronize()
    return (loss_a, d_a, m_a), (loss_b, d_b, m_b)


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("shape", [(3, 37, 12), (32, 883, 12)])
def test_static_loss_matches_existing_entry_point_twice(shape):
    from msgat_b200 import _lib
    lib = _lib.get()
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(20261105 + shape[1])
    for call in range(2):   # two calls in a row on fresh data: the second one meets the first one's scratch contents
        target = (torch.rand(shape, generator=g) * 300.0).to(dev)
        target[target < 20.0] = 0.0                      # zeros exercise the MAPE mask
        pred = (target + torch.randn(shape, generator=g).to(dev) * 60.0).contiguous()
        ref, new = _call_pair(lib, _lib, pred, target, 50.0, 0.0)
        for name, a, b in zip(("loss", "dpred", "metrics"), ref, new):
            assert torch.isfinite(b).all(), (call, name)
            assert _rel(b, a) < TOL, (call, name, _rel(b, a))
        # same kernel, same grid, same summation order: the results are in fact identical
        assert torch.equal(ref[0], new[0]) and torch.equal(ref[1], new[1]) and torch.equal(ref[2], new[2]), call


def test_static_loss_rejects_sizes_beyond_its_limit():
    from msgat_b200 import _lib
    lib = _lib.get()
    limit = int(lib.msgat_loss_static_max_n())
    assert limit == 1 << 28
    dev = torch.device("cuda")
    x = torch.zeros(16, device=dev)
    loss = torch.zeros(1, device=dev)
    rc = lib.msgat_loss_forward_static(_lib.ptr(x), _lib.ptr(x), C.c_int64(limit + 1), 1.0, 0.0, _lib.ptr(loss), None, None,
                                       _lib.stream_of(x))
    assert rc == -63
    assert b"limit" in lib.msgat_last_error()
    torch.cuda.synchronize()
