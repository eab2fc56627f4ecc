"""Build the drop-in module for a golden case and run it (shared by the emulated-kernel
tests on the build box and the GPU parity tests)."""
import torch

from msgat_b200 import models as M


def kind_of(case: str) -> str:
    for k in ("graph_attention", "temporal_attention", "channel_attention", "gacn", "tacn", "cacn", "meam", "tpc",
              "msgat48", "msgat72", "msgat96"):
        if case.startswith(k + "_"):
            return k
    raise KeyError(case)


def build_module(G):
    """Instantiate the drop-in class of golden case ``G`` and load the fixture's weights."""
    k, p, ins, meta = kind_of(G.case), G.params, G.inputs, G.meta
    if k.startswith("msgat"):
        X = ins["X"]
        mod = getattr(M, k)(n_components=X.shape[1], in_channels=X.shape[2], in_timesteps=12,
                            out_timesteps=p["tpcs.0.fc.bias"].shape[0], use_te=True, adj=p["adj"].float().clone())
    else:
        x = ins["signals"]
        C, N = x.shape[1], x.shape[2]
        if k == "graph_attention":
            mod = M.GraphAttention(C, 12)
        elif k == "temporal_attention":
            mod = M.TemporalAttention(C, N)
        elif k == "channel_attention":
            mod = M.ChannelAttention(N, 12)
        elif k == "gacn":
            mod = M.GACN(C, p["W"].shape[0], 12)
        elif k == "tacn":
            mod = M.TACN(C, p["seq.1.weight"].shape[0], N, meta["dilations"])
        elif k == "cacn":
            mod = M.CACN(C, p["seq.1.weight"].shape[0], N, 12)
        elif k == "meam":
            mod = M.MEAM(C, p["res.weight"].shape[0], N, 12, meta["dilations"])
        elif k == "tpc":
            chans = [C] + [p[f"tgacns.{j}.res.weight"].shape[0] for j in range(len(meta["dilations"]))]
            mod = M.TPC(chans, N, 12, p["fc.bias"].shape[0], meta["dilations"])
    missing = mod.load_state_dict({k2: v.float() for k2, v in p.items()}, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return mod


def run_case(G, device="cpu"):
    """-> (out, {param: grad}, {input: grad}) of the drop-in module on the golden inputs."""
    mod = build_module(G).to(device)
    names = [n for n in ("signals", "adjacency", "X", "H", "D") if n in G.inputs]
    args = []
    for n in names:
        v = G.inputs[n]
        v = v.float().to(device) if v.is_floating_point() else v.to(device)
        if n in ("signals", "X"):
            v.requires_grad_(True)
        args.append(v)
    out = mod(*args)
    (out * G.R.float().to(device)).sum().backward()
    grads = {n: prm.grad for n, prm in mod.named_parameters() if prm.grad is not None}
    gin = {n: a.grad for n, a in zip(names, args) if a.is_floating_point() and a.grad is not None}
    return out, grads, gin
