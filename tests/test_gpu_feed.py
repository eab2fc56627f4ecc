"""GPU: the device-resident sliding-window sampler (csrc/feed.cu, msgat_b200/data.py) against the batches the
reference loader produced (data_loader.py:68-115), and the double-buffered host feed."""
import os

import pytest
import torch

from feed_cases import FIXTURES, run_feed_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p) for p in FIXTURES])
def test_device_feed_matches_reference_loader(path):
    assert run_feed_case(path, torch.device("cuda")) >= 7


def test_window_gather_marks_out_of_range_windows():
    from msgat_b200 import data as D
    series = D.DeviceTimeSeries(torch.randn(2, 5, 100), 60, "cuda")
    X, H, Dd, Y = D.gather_windows(series, torch.tensor([5, 95]), [1], 12, 12)
    torch.cuda.synchronize()
    assert torch.isnan(X[0, ..., :7]).all() and not torch.isnan(X[0, ..., 7:]).any()     # window starts at t - 12 = -7
    assert not torch.isnan(X[1]).any()
    assert not torch.isnan(Y[0]).any() and torch.isnan(Y[1, :, 5:]).all() and not torch.isnan(Y[1, :, :5]).any()
    with pytest.raises(ValueError):
        D.WindowLoader(series, (5, 50), [1], 12, 12, 8)


def test_device_feeder_double_buffers_host_batches():
    """train.DeviceFeeder: batches staged from pinned host memory arrive intact and in order while the previous one is in use."""
    from msgat_b200.train import DeviceFeeder
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(3)
    host = [[torch.randn(4, 3, 1, 50, 12, generator=g).pin_memory(), torch.randint(0, 24, (4,), generator=g).pin_memory(),
             torch.randint(0, 7, (4,), generator=g).pin_memory(), torch.randn(4, 50, 12, generator=g).pin_memory()] for _ in range(5)]
    feeder = DeviceFeeder(host[0], dev)
    feeder.stage(host[0])
    sums = []
    for step in range(5):
        if step + 1 < 5:
            feeder.stage(host[step + 1])       # the next batch moves while this one is consumed
        i, (X, H, D, Y) = feeder.take()
        for a, b in zip((X, H, D, Y), host[step]):
            assert a.device.type == "cuda" and torch.equal(a.cpu(), b), step
        sums.append((X.sum() + Y.sum()).item())
        feeder.release(i)
    assert len(set(sums)) == 5
