"""CPU restatement of the reference's data feed (TEST INFRASTRUCTURE: only tests/ may import this).

Follows ``src/data_loader.py`` of the reference: ``normalize`` (:118-120), the split of the valid start indices
(:69-78), ``TimeSeriesSlice.__getitem__`` (:107-112) and the default collate of a batch.  numpy only.  Pinned by
``tests/golden/feed_*.npz``, minted by ``tests/golden/make_golden_feed.py`` from the unmodified reference loader.
"""
import numpy as np


def normalize(data_cnl, split):
    """data_loader.py:118-120 (torch.std_mean: unbiased)."""
    part = data_cnl[..., :split].astype(np.float64)   # statistics in fp64: an fp32 reduction depends on its order
    mean = part.mean(axis=-1, keepdims=True).astype(np.float32)
    std = part.std(axis=-1, keepdims=True, ddof=1).astype(np.float32)
    return ((data_cnl.astype(np.float32) - mean) / std).astype(np.float32)


def intervals(L, hours, tau, q):
    """data_loader.py:69-78."""
    in_timesteps = tau * max(hours)
    length = L - in_timesteps - q + 1
    s1, s2 = int(0.6 * length), int(0.8 * length)
    return [(in_timesteps, in_timesteps + s1), (in_timesteps + s1, in_timesteps + s2), (in_timesteps + s2, in_timesteps + length)]


def item(inputs_cnl, target_nl, t, hours, tau, q):
    """data_loader.py:107-112 for the absolute time index t (= i + interval[0])."""
    h = t // tau
    d = h // 24
    x = np.stack([inputs_cnl[..., t - hh * tau: t - hh * tau + tau] for hh in hours])
    y = target_nl[..., t: t + q]
    return x, h % 24, d % 7, y


def batch(inputs_cnl, target_nl, ts, hours, tau, q):
    """default collate of the items of a batch."""
    items = [item(inputs_cnl, target_nl, int(t), hours, tau, q) for t in ts]
    return (np.stack([i[0] for i in items]), np.array([i[1] for i in items], dtype=np.int64),
            np.array([i[2] for i in items], dtype=np.int64), np.stack([i[3] for i in items]))


def adjacency(edges, num_nodes):
    """data_loader.py:49-66: D^-1/2 (A + I) D^-1/2, symmetric 0/1 edges."""
    A = np.eye(num_nodes, dtype=np.float32)
    for s, d in edges:
        A[s, d] = A[d, s] = 1
    r = (1.0 / np.sqrt(A.sum(axis=1, dtype=np.float32))).astype(np.float32)
    return (r[:, None] * A) * r[None, :]
