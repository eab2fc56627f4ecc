"""Mint tests/golden/feed_*.npz from the UNMODIFIED reference loader (run in the build container only):

    python tests/golden/make_golden_feed.py

Writes a tiny synthetic dataset (meta.yaml, edge list, .npz series) into a temporary directory, runs the reference's
``DataLoaderForMSGAT`` (``/root/reference/src/data_loader.py``) on it from that directory and stores the raw series,
the edge list and what the reference produced: the adjacency, and for each of the three loaders the first two batches
and the last one (validation / evaluation are unshuffled, so their batch k starts at index k * batch_size of the
interval; the training batches are stored with the start indices recovered from H, D and Y).
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src"


def mint(tag, C, N, L, hours, tau, q, bs, seed):
    rng = np.random.default_rng(seed)
    data = (rng.standard_normal((L, N, C)) * 30 + 100 + np.arange(L)[:, None, None] * 0.37).astype(np.float32)
    edges = [(int(a), int(b)) for a, b in rng.integers(0, N, size=(2 * N, 2)) if a != b]
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "data"))
        np.savez(os.path.join(tmp, "data", "s.npz"), data=data)
        with open(os.path.join(tmp, "data", "s.csv"), "w") as f:
            f.write("from,to,cost\n")
            for a, b in edges:
                f.write(f"{a},{b},1.0\n")
        with open(os.path.join(tmp, "data", "meta.yaml"), "w") as f:
            f.write(f"S:\n  adj-file: data/s.csv\n  data-file: data/s.npz\n  num-nodes: {N}\n  num-channels: {C}\n"
                    f"  timesteps-per-hour: {tau}\n")
        os.chdir(tmp)
        try:
            sys.path.insert(0, REF)
            import data_loader as ref
            torch.manual_seed(seed)
            dl = ref.DataLoaderForMSGAT("S", hours, q, bs, 0)
            out = {"data": data, "edges": np.array(edges, dtype=np.int64), "hours": np.array(hours), "tau": tau, "q": q, "bs": bs,
                   "adj": dl.adj.numpy()}
            for name, loader in (("training", dl.training), ("validation", dl.validation), ("evaluation", dl.evaluation)):
                batches = list(loader)
                out[f"{name}_len"] = len(batches)
                out[f"{name}_interval"] = np.array(loader.dataset.interval)
                for k in sorted({0, 1, len(batches) - 1}):
                    x, h, d, y = batches[k]
                    out[f"{name}_{k}_X"], out[f"{name}_{k}_H"] = x.numpy(), h.numpy()
                    out[f"{name}_{k}_D"], out[f"{name}_{k}_Y"] = d.numpy(), y.numpy()
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(HERE, f"feed_{tag}.npz"), **out)
    print(tag, {k: getattr(v, "shape", v) for k, v in out.items() if not k.startswith(("training_", "validation_", "evaluation_"))})


if __name__ == "__main__":
    mint("C3N7_h123", C=3, N=7, L=150, hours=[1, 2, 3], tau=12, q=12, bs=8, seed=1)
    mint("C1N5_h1_24_168", C=1, N=5, L=168 * 12 + 700, hours=[1, 2, 24, 168], tau=12, q=12, bs=32, seed=2)
