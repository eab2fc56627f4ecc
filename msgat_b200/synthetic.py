"""Synthetic PEMS-shaped inputs.

The real datasets are Drive-hosted and unavailable offline, so every benchmark
and parity case uses tensors with the contract of the reference's loader
(``src/data_loader.py:106-112``): ``X [B,M,C,N,12]`` z-scored floats, ``H`` hour
and ``D`` weekday indices, ``Y [B,N,12]`` raw (un-normalised) targets, and an
adjacency normalised exactly like ``src/data_loader.py:59-66``
(``D^-1/2 (A+I) D^-1/2``).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Tuple

import numpy as np
import torch

# (N sensors, C channels, edges from the paper's dataset table, default -i list, batch)
PEMS_SHAPES = {
    "pemsd3": dict(N=358, C=1, E=547, M=4, B=64),
    "pemsd4": dict(N=307, C=3, E=340, M=5, B=64),
    "pemsd7": dict(N=883, C=1, E=866, M=5, B=32),
    "pemsd8": dict(N=170, C=3, E=295, M=5, B=64),
    "n4096": dict(N=4096, C=3, E=None, M=5, B=16),
}


def normalize_adjacency(A01: torch.Tensor) -> torch.Tensor:
    """Symmetric normalisation of a 0/1 adjacency that already has self loops."""
    d = A01.sum(1).rsqrt()
    return d[:, None] * A01 * d[None, :]


def road_graph(N: int, E: int, seed: int = 0, shuffle: bool = True) -> torch.Tensor:
    """Road-like graph: a spanning chain plus random short chords, ``E`` undirected
    edges in total; optionally relabelled by a random permutation (real PEMS node
    ids are not spatially sorted)."""
    rng = np.random.default_rng(seed)
    A = np.eye(N, dtype=np.float32)
    edges = set()
    for i in range(min(E, N - 1)):
        edges.add((i, i + 1))
    while len(edges) < E:
        s = int(rng.integers(0, N))
        hop = int(rng.integers(2, 12))
        d = min(N - 1, s + hop)
        if s != d:
            edges.add((min(s, d), max(s, d)))
    perm = rng.permutation(N) if shuffle else np.arange(N)
    for s, d in edges:
        A[perm[s], perm[d]] = A[perm[d], perm[s]] = 1.0
    return normalize_adjacency(torch.from_numpy(A))


def _morton_key(x: np.ndarray, y: np.ndarray, bits: int = 16) -> np.ndarray:
    xi = (x * ((1 << bits) - 1)).astype(np.uint64)
    yi = (y * ((1 << bits) - 1)).astype(np.uint64)
    key = np.zeros_like(xi)
    for b in range(bits):
        key |= ((xi >> np.uint64(b)) & np.uint64(1)) << np.uint64(2 * b)
        key |= ((yi >> np.uint64(b)) & np.uint64(1)) << np.uint64(2 * b + 1)
    return key


def geometric_graph(N: int, mean_degree: float = 8.0, seed: int = 0) -> torch.Tensor:
    """Distance-thresholded sensor graph on the unit square, nodes sorted along a
    Morton curve so that neighbours get nearby ids (BASELINE.json config 5)."""
    rng = np.random.default_rng(seed)
    pts = rng.random((N, 2))
    order = np.argsort(_morton_key(pts[:, 0], pts[:, 1]))
    pts = pts[order]
    r = math.sqrt(mean_degree / (math.pi * N))
    A = np.eye(N, dtype=np.float32)
    step = 1024
    for s in range(0, N, step):
        d2 = ((pts[s:s + step, None, :] - pts[None, :, :]) ** 2).sum(-1)
        A[s:s + step][d2 < r * r] = 1.0
    A = np.maximum(A, A.T)
    return normalize_adjacency(torch.from_numpy(A))


def make_adjacency(name: str, seed: int = 0, shuffle: bool = True) -> torch.Tensor:
    s = PEMS_SHAPES[name]
    if s["E"] is None:
        return geometric_graph(s["N"], 8.0, seed)
    return road_graph(s["N"], s["E"], seed, shuffle)


@dataclass
class Batch:
    X: torch.Tensor  # [B, M, C, N, 12] f32
    H: torch.Tensor  # [B] int64 in [0, 24)
    D: torch.Tensor  # [B] int64 in [0, 7)
    Y: torch.Tensor  # [B, N, Tout] f32


def make_batch(B: int, M: int, C: int, N: int, T: int = 12, Tout: int = 12, seed: int = 1234,
               pin: bool = False) -> Batch:
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(B, M, C, N, T, generator=g)
    H = torch.randint(0, 24, (B,), generator=g)
    D = torch.randint(0, 7, (B,), generator=g)
    Y = (200.0 + 100.0 * torch.randn(B, N, Tout, generator=g)).clamp_min(0.0)
    if pin and torch.cuda.is_available():
        X, H, D, Y = X.pin_memory(), H.pin_memory(), D.pin_memory(), Y.pin_memory()
    return Batch(X, H, D, Y)


def shard(batch: Batch, rank: int, world: int) -> Batch:
    """Rank r of R owns samples [r*B/R, (r+1)*B/R)."""
    B = batch.X.shape[0]
    assert B % world == 0, "global batch must divide evenly across ranks"
    lo, hi = rank * B // world, (rank + 1) * B // world
    return Batch(batch.X[lo:hi], batch.H[lo:hi], batch.D[lo:hi], batch.Y[lo:hi])


def tile_occupancy(adj: torch.Tensor, tile: int = 128) -> Tuple[int, int]:
    """(non-empty tiles, total tiles) of the adjacency at ``tile`` x ``tile``."""
    N = adj.shape[0]
    nt = (N + tile - 1) // tile
    pad = nt * tile - N
    a = torch.nn.functional.pad(adj != 0, (0, pad, 0, pad))
    occ = a.view(nt, tile, nt, tile).any(3).any(1)
    return int(occ.sum()), nt * nt
