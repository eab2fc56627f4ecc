import pytest
import torch

from head_cases import HEAD_CASES, run_head_case


@pytest.mark.gpu
@pytest.mark.parametrize("case", HEAD_CASES)
def test_head_matches_fp64_reference(case):
    """fp32 path, tolerance 1e-5 relative L2 per tensor (BASELINE.json north_star), forward and every gradient."""
    errs = run_head_case(case, torch.device("cuda"))
    bad = {k: v for k, v in errs.items() if not v < 1e-5}
    assert not bad, f"{case}: {bad}"
