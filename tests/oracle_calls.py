"""Dispatch a golden case name to the matching oracle function (test helper)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import msgat_oracle as O  # noqa: E402


def kind_of(case: str) -> str:
    for k in ("graph_attention", "temporal_attention", "channel_attention", "gacn", "tacn", "cacn", "meam", "tpc",
              "msgat48", "msgat72", "msgat96"):
        if case.startswith(k + "_"):
            return k
    raise KeyError(case)


def oracle_forward(case, meta, p, ins):
    """p: dict name->tensor (reference state_dict names), ins: dict of inputs."""
    k = kind_of(case)
    if k == "graph_attention":
        return O.graph_attention(ins["signals"], ins["adjacency"], p["Wg"], p["alpha"])
    if k == "temporal_attention":
        return O.temporal_attention(ins["signals"], p["Wt1"], p["Wt2"], p["alpha"])
    if k == "channel_attention":
        return O.channel_attention(ins["signals"], p["Wc"], p["alpha"])
    if k == "gacn":
        return O.gacn(ins["signals"], ins["adjacency"], p, "")
    if k == "tacn":
        return O.tacn(ins["signals"], p, "", meta["dilations"])
    if k == "cacn":
        return O.cacn(ins["signals"], p, "")
    if k == "meam":
        return O.meam(ins["signals"], ins["adjacency"], p, "", meta["dilations"])
    if k == "tpc":
        return O.tpc(ins["signals"], ins["adjacency"], p, "", meta["dilations"])
    return O.msgat(ins["X"], ins["H"], ins["D"], p, O.MODEL_DILATIONS[k])
