"""Mint golden vectors from the UNMODIFIED reference (run in the build container).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Imports ``models`` from ``/root/reference/src`` (read-only), builds each layer
class of the hot path with the reference's own ``reset_parameters``-style init,
runs forward and backward in fp64 (``copy.deepcopy(m).double()``) and in fp32
as shipped, and stores inputs, weights, outputs and every gradient in
``tests/golden/<case>.npz``.  The loss is ``sum(out * R)`` with a fixed random
``R`` so that all gradients are non-trivial.  ``/root/reference`` does not exist on
the GPU box; only the .npz files travel.
"""
from __future__ import annotations

import copy
import json
import os
import sys
import zlib

import numpy as np
import torch

REF = os.environ.get("MSGAT_REFERENCE", "/root/reference/src")
sys.path.insert(0, REF)
sys.dont_write_bytecode = True
from models import attention as ref_att  # noqa: E402
from models import msgat as ref_msgat  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from msgat_b200.synthetic import road_graph  # noqa: E402


def init_like_reference(module: torch.nn.Module, scale: dict | None = None):
    """MSGAT.reset_parameters rule (msgat.py:206-217) applied to a sub-module."""
    for name, prm in module.named_parameters():
        if not prm.requires_grad:
            continue
        if prm.ndim >= 2:
            torch.nn.init.xavier_normal_(prm)
        else:
            s = prm.size(0) ** -0.5
            torch.nn.init.uniform_(prm, -s, s)
        if scale:
            for key, f in scale.items():
                if name.endswith(key):
                    prm.data.mul_(f)


def run(module, inputs, R, dtype):
    m = copy.deepcopy(module).to(dtype)
    ins = []
    for v in inputs:
        if v.is_floating_point():
            v = v.detach().to(dtype).clone()
            v.requires_grad_(True)
        ins.append(v)
    out = m(*ins)
    (out * R.to(dtype)).sum().backward()
    grads = {n: p.grad.detach().double().numpy() for n, p in m.named_parameters() if p.grad is not None}
    gins = [v.grad.detach().double().numpy() if v.is_floating_point() and v.grad is not None else None for v in ins]
    return out.detach().double().numpy(), grads, gins


def rel(a, b):
    return np.float64(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def record(case, module, inputs, input_names, meta):
    torch.manual_seed(zlib.crc32(case.encode()) % 100000)
    with torch.no_grad():
        probe = module(*inputs)
    R = torch.randn(probe.shape, dtype=torch.float32).double()  # fp32-representable
    out64, g64, gin64 = run(module, inputs, R, torch.float64)
    out32, g32, gin32 = run(module, inputs, R, torch.float32)
    # the fp32 run of the reference is kept only as its rel-L2 distance to the fp64 run:
    # that is the reference's own noise floor for each tensor
    # large arrays (out, input grads, R) are stored rounded to fp32 (rel. error 6e-8, far below
    # the 1e-5 gate) to keep the fixtures small; parameter grads stay fp64
    blob = {"R": R.numpy().astype(np.float32), "out64": out64.astype(np.float32), "err32.out": rel(out32, out64),
            "meta": np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)}
    for n, v in zip(input_names, inputs):
        blob["in." + n] = v.detach().numpy()
    for n, prm in module.state_dict().items():
        blob["p." + n] = prm.detach().numpy()
    for n, v in g64.items():
        blob["g64." + n] = v
        blob["err32.g." + n] = rel(g32[n], v)
    for n, a, b in zip(input_names, gin64, gin32):
        if a is not None:
            blob["gin64." + n] = a.astype(np.float32)
            blob["err32.gin." + n] = rel(b, a)
    np.savez_compressed(os.path.join(HERE, case + ".npz"), **blob)
    print(f"{case:28s} out {tuple(out64.shape)}  params {len(g64)}")


def adjacency(N, kind, seed):
    if kind == "road":
        return road_graph(N, int(1.3 * N), seed=seed)
    g = torch.Generator().manual_seed(seed)
    if kind == "dense_signed":  # arbitrary real weights: the mask is a plain multiply
        return torch.randn(N, N, generator=g) * (torch.rand(N, N, generator=g) < 0.5)
    if kind == "empty_rows":
        A = road_graph(N, int(1.3 * N), seed=seed)
        A[3] = 0
        A[N - 1] = 0
        A[:, 5] = 0
        return A
    raise ValueError(kind)


def main():
    B, T = 2, 12
    PEAK = {"Wg": 30.0, "Wc": 30.0, "Wt1": 5.0, "Wt2": 5.0}
    for tag, scale in (("", None), ("_peaked", PEAK)):
        for N, C, kind in ((23, 5, "road"), (37, 1, "dense_signed"), (19, 3, "empty_rows")):
            adj = adjacency(N, kind, seed=N)
            torch.manual_seed(N * 100 + C)
            x = torch.randn(B, C, N, T)
            sfx = f"_N{N}C{C}{tag}"
            m = ref_att.GraphAttention(C, T); init_like_reference(m, scale)
            record("graph_attention" + sfx, m, [x, adj], ["signals", "adjacency"], dict(C=C, N=N, T=T, adj=kind))
            m = ref_att.TemporalAttention(C, N); init_like_reference(m, scale)
            record("temporal_attention" + sfx, m, [x], ["signals"], dict(C=C, N=N, T=T))
            m = ref_att.ChannelAttention(N, T); init_like_reference(m, scale)
            record("channel_attention" + sfx, m, [x], ["signals"], dict(C=C, N=N, T=T))
            m = ref_msgat.GACN(C, 8, T); init_like_reference(m, scale)
            record("gacn" + sfx, m, [x, adj], ["signals", "adjacency"], dict(C=C, O=8, N=N, T=T, adj=kind))
            for dil in ([1, 2], [2, 4], [1, 1, 2, 2]) if scale else ([1, 2],):
                m = ref_msgat.TACN(C, 8, N, dil); init_like_reference(m, scale)
                record(f"tacn{sfx}_d{''.join(map(str, dil))}", m, [x], ["signals"], dict(C=C, O=8, N=N, T=T, dilations=dil))
            m = ref_msgat.CACN(C, 8, N, T); init_like_reference(m, scale)
            record("cacn" + sfx, m, [x], ["signals"], dict(C=C, O=8, N=N, T=T))
            for dil in ([4, 4],) if scale else ([1, 2],):
                m = ref_msgat.MEAM(C, 24, N, T, dil); init_like_reference(m, scale)
                record(f"meam{sfx}_d{''.join(map(str, dil))}", m, [x, adj], ["signals", "adjacency"],
                       dict(C=C, O=24, N=N, T=T, dilations=dil, adj=kind))
    # a wide MEAM with the real layer-2 shape (72 -> 72) on a small graph
    adj = adjacency(29, "road", 5)
    torch.manual_seed(11)
    x = torch.randn(B, 72, 29, T)
    m = ref_msgat.MEAM(72, 72, 29, T, [2, 4]); init_like_reference(m, PEAK)
    record("meam_N29C72_peaked_d24", m, [x, adj], ["signals", "adjacency"], dict(C=72, O=72, N=29, T=T, dilations=[2, 4], adj="road"))
    # TPC and full model (small widths through the MSGAT ctor, real factories at tiny N)
    for N, C in ((23, 3), (31, 1)):
        adj = adjacency(N, "road", N)
        torch.manual_seed(N)
        x = torch.randn(B, C, N, T)
        m = ref_msgat.TPC([C, 24, 24], N, T, 12, [[1, 2], [2, 4]]); init_like_reference(m)
        record(f"tpc_N{N}C{C}", m, [x, adj], ["signals", "adjacency"], dict(channels=[C, 24, 24], N=N, T=T, Tout=12, dilations=[[1, 2], [2, 4]]))
    for name, factory, N, C, M in (("msgat72", ref_msgat.msgat72, 21, 3, 2), ("msgat48", ref_msgat.msgat48, 17, 1, 3), ("msgat96", ref_msgat.msgat96, 13, 3, 2)):
        adj = adjacency(N, "road", N)
        torch.manual_seed(N + 1)
        m = factory(M, C, in_timesteps=T, out_timesteps=12, use_te=True, adj=adj)
        X = torch.randn(B + 1, M, C, N, T)
        H = torch.randint(0, 24, (B + 1,))
        D = torch.randint(0, 7, (B + 1,))
        record(f"{name}_N{N}C{C}M{M}", m, [X, H, D], ["X", "H", "D"], dict(model=name, M=M, C=C, N=N, T=T, Tout=12))
    # parameter names / shapes / order of the real factories (the drop-in contract)
    contract = {}
    for name, factory in (("msgat48", ref_msgat.msgat48), ("msgat72", ref_msgat.msgat72), ("msgat96", ref_msgat.msgat96)):
        for C, M in ((1, 4), (3, 5)):
            m = factory(M, C, in_timesteps=12, out_timesteps=12, use_te=True, adj=torch.eye(11))
            contract[f"{name}_C{C}_M{M}"] = dict(
                state_dict=[[k, list(v.shape)] for k, v in m.state_dict().items()],
                parameters=[[k, list(v.shape), bool(v.requires_grad)] for k, v in m.named_parameters()])
    with open(os.path.join(HERE, "state_dict_contract.json"), "w") as f:
        json.dump(contract, f)
    print("wrote state_dict_contract.json")


if __name__ == "__main__":
    main()
