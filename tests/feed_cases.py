"""Shared cases of the data feed: golden fixtures minted from the reference loader (tests/golden/make_golden_feed.py)."""
import glob
import os
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "feed_*.npz")))


def load(path):
    z = np.load(path)
    return {k: z[k] for k in z.files}


def stored_batches(g, name):
    ks = sorted({0, 1, int(g[f"{name}_len"]) - 1})
    return [(k, g[f"{name}_{k}_X"], g[f"{name}_{k}_H"], g[f"{name}_{k}_D"], g[f"{name}_{k}_Y"]) for k in ks]


def recover_indices(g, Y, interval):
    """start index of every sample of a (shuffled) batch: the raw target has a trend, so its window is unique"""
    target = g["data"].transpose(2, 1, 0)[0]   # [N, L]
    q = int(g["q"])
    ts = []
    for b in range(Y.shape[0]):
        hits = [t for t in range(int(interval[0]), int(interval[1])) if np.array_equal(target[:, t:t + q], Y[b])]
        assert len(hits) == 1, hits
        ts.append(hits[0])
    return np.array(ts, dtype=np.int64)


def write_dataset(g, tmp):
    """the fixture's series and edge list as the files the reference reads (data/meta.yaml layout)"""
    os.makedirs(os.path.join(tmp, "data"), exist_ok=True)
    np.savez(os.path.join(tmp, "data", "s.npz"), data=g["data"])
    with open(os.path.join(tmp, "data", "s.csv"), "w") as f:
        f.write("from,to,cost\n")
        for a, b in g["edges"]:
            f.write(f"{int(a)},{int(b)},1.0\n")
    L, N, C = g["data"].shape
    meta = os.path.join(tmp, "data", "meta.yaml")
    with open(meta, "w") as f:
        f.write(f"S:\n  adj-file: {tmp}/data/s.csv\n  data-file: {tmp}/data/s.npz\n  num-nodes: {N}\n  num-channels: {C}\n"
                f"  timesteps-per-hour: {int(g['tau'])}\n")
    return meta


def run_feed_case(path, device):
    """msgat_b200.data against what the reference loader produced: bit-exact X / H / D / Y, adjacency to 1 ulp."""
    from msgat_b200 import data as D
    g = load(path)
    hours, tau, q, bs = [int(h) for h in g["hours"]], int(g["tau"]), int(g["q"]), int(g["bs"])
    with tempfile.TemporaryDirectory() as tmp:
        meta = write_dataset(g, tmp)
        dl = D.DeviceDataLoaderForMSGAT("S", hours, q, bs, meta_file=meta, device=device, seed=0)
    assert np.allclose(dl.adj.numpy(), g["adj"], rtol=2e-7, atol=0)
    checked = 0
    for name, loader in (("training", dl.training), ("validation", dl.validation), ("evaluation", dl.evaluation)):
        assert len(loader) == int(g[f"{name}_len"]), name
        assert tuple(loader.interval) == tuple(int(v) for v in g[f"{name}_interval"]), name
        if name != "training":
            mine = list(loader)
            assert len(mine) == len(loader)
        for k, X, H, Dd, Y in stored_batches(g, name):
            if name == "training":   # shuffled: replay the reference's sample order
                t = torch.from_numpy(recover_indices(g, Y, g[f"{name}_interval"]))
                got = D.gather_windows(dl.series, t.to(device), hours, tau, q)
            else:
                got = mine[k]
            for a, b, nm in zip(got, (X, H, Dd, Y), "XHDY"):
                assert tuple(a.shape) == b.shape, (name, k, nm, a.shape, b.shape)
                assert np.array_equal(a.cpu().numpy(), b), (name, k, nm)
            checked += 1
    # a training epoch visits every start index of its interval exactly once
    lo, hi = dl.training.interval
    seen = torch.cat([dl.training.indices()]).sort().values
    assert torch.equal(seen, torch.arange(lo, hi))
    return checked
