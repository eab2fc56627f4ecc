"""CPU: the drop-in modules keep the reference's state_dict names, shapes and ORDER and its parameter
registration order (`model.parameters()` order drives Adam state in checkpoints, engine.py:106,140)."""
import json
import os

import torch


def _contract():
    with open(os.path.join(os.path.dirname(__file__), "golden", "state_dict_contract.json")) as f:
        return json.load(f)


def test_state_dict_contract():
    """msgat_b200.models against tests/golden/state_dict_contract.json (minted from the unmodified reference by
    tests/golden/make_golden.py): all six factory/shape combinations."""
    from msgat_b200 import models
    for key, c in _contract().items():
        name, C, M = key.split("_")
        model = getattr(models, name)(n_components=int(M[1:]), in_channels=int(C[1:]), in_timesteps=12,
                                      out_timesteps=12, use_te=True, adj=torch.eye(11))
        got_sd = [[k, list(v.shape)] for k, v in model.state_dict().items()]
        assert got_sd == c["state_dict"], key
        got_p = [[k, list(v.shape), bool(v.requires_grad)] for k, v in model.named_parameters()]
        assert got_p == c["parameters"], key


def test_static_gate_parameters_without_time_embedding():
    """use_te=False registers the static gate W [M, N, Tout] instead of the embeddings (msgat.py:186-189)."""
    from msgat_b200 import models
    model = models.msgat72(n_components=3, in_channels=1, in_timesteps=12, out_timesteps=12, use_te=False,
                           adj=torch.eye(7))
    names = [n for n, _ in model.named_parameters()]
    assert names[0] == "W" and tuple(model.W.shape) == (3, 7, 12) and names[1] == "adj"
    assert not any(n.startswith("te.") for n in names)
