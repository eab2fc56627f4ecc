"""Data feed with the time series resident on the GPU.

Same protocol as the reference's ``DataLoaderForMSGAT`` (``src/data_loader.py:15-88``): ``.adj`` and three loaders
``.training`` / ``.validation`` / ``.evaluation`` that yield ``(X[B,M,C,N,tau], H[B], D[B], Y[B,N,q])`` per batch,
60 / 20 / 20 % split over the valid start indices, inputs z-scored with the statistics of the training part, the
target left raw (channel 0), shuffle on the training loader only, last batch kept short.  The difference is where
the work happens: the reference slices and stacks M windows per SAMPLE in Python (``data_loader.py:107-112``), collates
on the host and copies 8 MB per PEMSD7 batch to the GPU; here a batch is B time indices and one gather kernel
(``csrc/feed.cu``) cuts every window out of the series in HBM.  The files are the reference's (``data/meta.yaml``, the
``.npz`` series with key ``data`` of shape ``[L, N, C]``, the edge-list ``.csv`` with one header line).
"""
from __future__ import annotations

import ctypes as C
from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib


def load_adjacency_csv(path: str, num_nodes: int) -> torch.Tensor:
    """``D^-1/2 (A + I) D^-1/2`` from an edge list ``src,dst,cost`` with one header line (data_loader.py:49-66)."""
    A = torch.eye(num_nodes)
    with open(path, "r") as f:
        for line in f.readlines()[1:]:
            if not line.strip():
                continue
            src, dst = line.split(",")[:2]
            A[int(src), int(dst)] = A[int(dst), int(src)] = 1
    d = A.sum(dim=1).rsqrt()
    return d[:, None] * A * d[None, :]


def normalize(tensor: torch.Tensor, split: int) -> torch.Tensor:
    """z-score along time with the statistics of ``[..., :split]`` (data_loader.py:118-120; unbiased std)."""
    std, mean = torch.std_mean(tensor[..., :split], dim=-1, keepdim=True)
    return (tensor - mean) / std


class DeviceTimeSeries:
    """The normalised series ``inputs [C,N,L]`` and the raw target ``[N,L]`` (channel 0), resident on ``device``."""

    def __init__(self, data_cnl: torch.Tensor, split: int, device):
        data = data_cnl.float()
        self.inputs = normalize(data, split).contiguous().to(device)     # data_loader.py:80
        self.target = data[0].contiguous().to(device)                    # data_loader.py:83 (data[0])
        self.C, self.N, self.L = (int(v) for v in data.shape)
        self.device = torch.device(device)

    @staticmethod
    def from_npz(path: str, split: int, device) -> "DeviceTimeSeries":
        data = torch.from_numpy(np.load(path)["data"]).float().transpose(0, -1)   # [L,N,C] -> [C,N,L], data_loader.py:71
        return DeviceTimeSeries(data, split, device)


def gather_windows(series: DeviceTimeSeries, t: torch.Tensor, hours: Sequence[int], tau: int, q: int,
                   out: Optional[Tuple[torch.Tensor, ...]] = None):
    """One batch of the sliding-window sampler: ``t`` is an int64 device tensor of start indices."""
    lib = _lib.get()
    t = t.to(device=series.inputs.device, dtype=torch.int64).contiguous()
    B, M = int(t.numel()), len(hours)
    dev = series.inputs.device
    if out is None:
        X = torch.empty(B, M, series.C, series.N, tau, device=dev)
        H = torch.empty(B, dtype=torch.int64, device=dev)
        D = torch.empty(B, dtype=torch.int64, device=dev)
        Y = torch.empty(B, series.N, q, device=dev)
    else:
        X, H, D, Y = out
    hrs = (C.c_int32 * M)(*[int(h) for h in hours])
    _lib.check(lib.msgat_window_gather(_lib.ptr(series.inputs), _lib.ptr(series.target), _lib.ptr(t), B, series.C, series.N,
                                       series.L, hrs, M, int(tau), int(q), _lib.ptr(X), _lib.ptr(H), _lib.ptr(D), _lib.ptr(Y),
                                       _lib.stream_of(series.inputs)), "msgat_window_gather")
    return X, H, D, Y


class WindowLoader:
    """Iterates an interval of start indices in batches (the ``DataLoader(TimeSeriesSlice(...))`` of data_loader.py:81-88)."""

    def __init__(self, series: DeviceTimeSeries, interval: Tuple[int, int], hours: Sequence[int], out_timesteps: int,
                 timesteps_per_hour: int, batch_size: int, shuffle: bool = False, seed: Optional[int] = None):
        lo, hi = int(interval[0]), int(interval[1])
        tau, q = int(timesteps_per_hour), int(out_timesteps)
        if lo - max(hours) * tau < 0 or hi - 1 + q > series.L:
            raise ValueError(f"interval {interval} leaves the series of {series.L} steps (hours {list(hours)}, tau {tau}, q {q})")
        if batch_size > 256:
            raise ValueError("batch_size > 256 is not supported by msgat_window_gather")
        self.series, self.interval, self.hours, self.q, self.tau = series, (lo, hi), list(hours), q, tau
        self.batch_size, self.shuffle = int(batch_size), shuffle
        self.generator = torch.Generator()
        if seed is not None:
            self.generator.manual_seed(seed)

    def __len__(self) -> int:
        n = self.interval[1] - self.interval[0]
        return (n + self.batch_size - 1) // self.batch_size

    def indices(self) -> torch.Tensor:
        n = self.interval[1] - self.interval[0]
        order = torch.randperm(n, generator=self.generator) if self.shuffle else torch.arange(n)
        return order + self.interval[0]

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]]:
        idx = self.indices().to(self.series.device)
        for s in range(0, idx.numel(), self.batch_size):
            yield gather_windows(self.series, idx[s:s + self.batch_size], self.hours, self.tau, self.q)


class DeviceDataLoaderForMSGAT:
    """``DataLoaderForMSGAT`` (data_loader.py:15-88) with the series and the sampling on the GPU."""

    def __init__(self, name: str, in_hours: List[int], out_timesteps: int, batch_size: int, num_workers: int = 0,
                 meta_file: str = "data/meta.yaml", device="cuda", seed: Optional[int] = None):
        import yaml
        with open(meta_file, "r") as f:
            meta = yaml.safe_load(f)[name]
        self.adj_file, self.data_file = meta["adj-file"], meta["data-file"]
        self.num_nodes, self.num_channels = meta["num-nodes"], meta["num-channels"]
        self.timesteps_per_hour = meta["timesteps-per-hour"]
        self.in_hours, self.out_timesteps, self.batch_size = list(in_hours), out_timesteps, batch_size
        self.adj = load_adjacency_csv(self.adj_file, self.num_nodes)
        tau = self.timesteps_per_hour
        in_timesteps = tau * max(self.in_hours)
        data = torch.from_numpy(np.load(self.data_file)["data"]).float().transpose(0, -1)
        length = data.size(-1) - in_timesteps - out_timesteps + 1
        split1, split2 = int(0.6 * length), int(0.8 * length)
        self.intervals = [(in_timesteps, in_timesteps + split1), (in_timesteps + split1, in_timesteps + split2),
                          (in_timesteps + split2, in_timesteps + length)]                      # data_loader.py:74-78
        self.series = DeviceTimeSeries(data, in_timesteps + split1, device)
        self.training, self.validation, self.evaluation = [
            WindowLoader(self.series, iv, self.in_hours, out_timesteps, tau, batch_size, shuffle=(i == 0), seed=seed)
            for i, iv in enumerate(self.intervals)]
