"""CPU: the numpy restatement of the reference's data feed (oracle/feed_oracle.py) against the golden fixtures minted
from the unmodified reference loader, and the device feed's host logic through the host-fiber emulator."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import emu_backend  # noqa: E402
from feed_cases import FIXTURES, load, recover_indices, run_feed_case, stored_batches  # noqa: E402
from oracle import feed_oracle as F  # noqa: E402


@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p) for p in FIXTURES])
def test_feed_oracle_matches_reference_loader(path):
    g = load(path)
    hours, tau, q, bs = [int(h) for h in g["hours"]], int(g["tau"]), int(g["q"]), int(g["bs"])
    data = g["data"].transpose(2, 1, 0).astype(np.float32)   # [C, N, L]
    iv = F.intervals(data.shape[-1], hours, tau, q)
    inputs = F.normalize(data, iv[0][1])
    assert np.allclose(F.adjacency(g["edges"], data.shape[1]), g["adj"], rtol=2e-7, atol=0)
    for name, interval in zip(("training", "validation", "evaluation"), iv):
        assert tuple(g[f"{name}_interval"]) == interval
        n = interval[1] - interval[0]
        assert int(g[f"{name}_len"]) == (n + bs - 1) // bs
        for k, X, H, D, Y in stored_batches(g, name):
            ts = recover_indices(g, Y, interval) if name == "training" else np.arange(interval[0] + k * bs, min(interval[0] + (k + 1) * bs, interval[1]))
            x, h, d, y = F.batch(inputs, data[0], ts, hours, tau, q)
            assert np.array_equal(h, H) and np.array_equal(d, D) and np.array_equal(y, Y)
            # the z-score statistics are fp32 reductions in the reference (order-dependent): 1e-5 of a unit-variance value
            assert x.shape == X.shape and np.allclose(x, X, rtol=0, atol=1e-5)


@pytest.fixture()
def emulated_kernels():
    from msgat_b200 import _lib
    saved = (_lib._LIB, _lib._DEBUG_BACKEND)
    emu_backend.install()
    yield
    _lib._LIB, _lib._DEBUG_BACKEND = saved


@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p) for p in FIXTURES])
def test_emulated_device_feed_matches_reference_loader(emulated_kernels, path):
    import torch
    assert run_feed_case(path, torch.device("cpu")) >= 7
