"""Kernel LOGIC check on the GPU-less build box: the same .cu sources compiled for the host
fiber emulator (tests/emu) run the drop-in modules on the golden cases minted from the
reference.  This debugs indexing and math before GPU minutes are spent; the parity claim
itself is made by tests/test_gpu_parity.py on the real device."""
import pytest

import emu_backend
from conftest import Golden
from layer_cases import run_case
from parity import check_case

CASES = ["graph_attention_N23C5_peaked", "temporal_attention_N19C3_peaked", "channel_attention_N23C5",
         "gacn_N37C1_peaked", "gacn_N23C5", "tacn_N23C5_peaked_d1122", "tacn_N19C3_d12", "cacn_N19C3_peaked",
         "meam_N37C1_d12", "meam_N23C5_peaked_d44", "meam_N29C72_peaked_d24", "tpc_N31C1", "msgat72_N21C3M2"]


@pytest.fixture(scope="module", autouse=True)
def emulated_kernels():
    from msgat_b200 import _lib
    saved = (_lib._LIB, _lib._DEBUG_BACKEND)
    emu_backend.install()
    yield
    _lib._LIB, _lib._DEBUG_BACKEND = saved


@pytest.mark.parametrize("case", CASES)
def test_emulated_kernels_match_reference_golden(case):
    G = Golden(case)
    out, grads, gin = run_case(G)
    fails, _ = check_case(G, out, grads, gin)
    assert not fails, "\n".join(fails)
