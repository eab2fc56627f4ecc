"""GPU: device-side Huber loss + streaming metrics + flat Adam against the oracle (loss.py:51-52,
metrics.py:20-35, engine.py:106)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_loss_and_metrics_match_oracle():
    from loss_cases import run_loss_case
    for shape in [(3, 37, 12), (32, 883, 12)]:
        errs = run_loss_case(torch.device("cuda"), n_shape=shape)
        bad = {k: v for k, v in errs.items() if not v < 2e-6}
        assert not bad, (shape, errs)


def test_flat_adam_matches_torch_adam():
    from loss_cases import run_adam_case
    assert run_adam_case(torch.device("cuda")) < 1e-5
