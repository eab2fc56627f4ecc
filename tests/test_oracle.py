"""Pin the CPU oracle against the fixtures minted from the reference itself.

Every golden case holds the reference's fp64 forward output, input gradient and
all parameter gradients for loss = sum(out * R).  The oracle, run in fp64 on
the same inputs and weights, must reproduce them to rounding (the fixtures
store the large arrays rounded to fp32, hence 3e-7 there and 1e-10 on the
parameter gradients, which are stored in fp64).
"""
import pytest
import torch

from conftest import Golden, golden_cases, rel_l2
from oracle_calls import oracle_forward

ALL = golden_cases("")


@pytest.mark.parametrize("case", ALL)
def test_oracle_matches_reference_golden(case):
    G = Golden(case)
    p = {k: v.double().requires_grad_(v.is_floating_point() and k != "adj") for k, v in G.params.items()}
    ins = {}
    for k, v in G.inputs.items():
        ins[k] = v.double().requires_grad_(True) if v.is_floating_point() else v
    out = oracle_forward(case, G.meta, p, ins)
    assert out.shape == G.out.shape
    assert rel_l2(out, G.out) < 3e-7, "forward"
    (out * G.R.double()).sum().backward()
    for name, g in G.grads.items():
        assert p[name].grad is not None, name
        assert rel_l2(p[name].grad, g) < 1e-10, f"grad {name}"
    for name, g in G.gin.items():
        if name == "adjacency":
            continue
        assert rel_l2(ins[name].grad, g) < 3e-7, f"input grad {name}"


def test_oracle_fp32_noise_is_at_reference_level():
    """The oracle in fp32 sits at the same distance from fp64 as the reference's
    own fp32 run (same ATen ops), so it is a fair CPU baseline."""
    G = Golden("meam_N29C72_peaked_d24")
    p = {k: v.float() for k, v in G.params.items()}
    ins = {k: v.float() for k, v in G.inputs.items()}
    out = oracle_forward(G.case, G.meta, p, ins)
    assert rel_l2(out, G.out) < max(1e-5, 10 * G.noise("out"))


def test_init_params_matches_contract():
    import json
    import os
    from oracle import msgat_oracle as O
    with open(os.path.join(os.path.dirname(__file__), "golden", "state_dict_contract.json")) as f:
        contract = json.load(f)
    for key, c in contract.items():
        model, C, M = key.split("_")
        p = O.init_params(model, int(M[1:]), int(C[1:]), torch.eye(11))
        want = {k: tuple(s) for k, s in c["state_dict"]}
        got = {k: tuple(v.shape) for k, v in p.items()}
        assert got == want, key
