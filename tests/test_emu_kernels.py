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


@pytest.mark.parametrize("nt", ["16", "32", "64"])
def test_emulated_fix_pass_tile_sizes(nt, monkeypatch):
    """k_bwd_fix picks 64- / 32-node tiles for narrow inputs; MSGAT_FIX_NT forces a size.  Every size must give the golden
    gradients on a first-layer (C = 1) and a C = 5 operator (the partial-sum layout follows the tile count)."""
    monkeypatch.setenv("MSGAT_FIX_NT", nt)
    for case in ("meam_N37C1_d12", "meam_N23C5_peaked_d44"):
        G = Golden(case)
        out, grads, gin = run_case(G)
        fails, _ = check_case(G, out, grads, gin)
        assert not fails, f"MSGAT_FIX_NT={nt} {case}:\n" + "\n".join(fails)


def test_emulated_head_kernels_ragged_shapes():
    """Streaming (Tout = 12) and generic head kernels on ragged node / channel counts (tests/head_cases.py)."""
    import torch
    from head_cases import HEAD_CASES, run_head_case
    for case in HEAD_CASES:
        errs = run_head_case(case, torch.device("cpu"))
        bad = {k: v for k, v in errs.items() if not v < 1e-5}
        assert not bad, f"{case}: {bad}"


@pytest.mark.parametrize("N", [37, 300])
def test_emulated_tensor_core_score_passes(N):
    """csrc/spatial_tc.cu (TF32 mode: the N x N score passes of attention.py:34-36 on tcgen05, forward and backward)
    through the emulated tcgen05 / TMEM model: one and three column tiles, ragged last tile, against the fp64 oracle
    at the TF32 bar."""
    import torch
    from msgat_b200 import _lib, models, synthetic
    from oracle import msgat_oracle as O
    torch.manual_seed(N)
    adj = synthetic.geometric_graph(N, mean_degree=5.0, seed=1)
    mod = models.GACN(3, 8, 12)
    for p in mod.parameters():
        torch.nn.init.normal_(p, std=0.5) if p.dim() < 2 else torch.nn.init.xavier_normal_(p)
    mod.precision = mod.gatt.precision = _lib.PREC_TF32
    g = torch.Generator().manual_seed(N + 1)
    x = torch.randn(2, 3, N, 12, generator=g)
    R = torch.randn(2, 8, N, 12, generator=g)
    xr = x.clone().requires_grad_(True)
    out = mod(xr, adj)
    (out * R).sum().backward()
    p = {k: v.detach().clone().double().requires_grad_(True) for k, v in mod.state_dict().items()}
    xo = x.clone().double().requires_grad_(True)
    ref = O.gacn(xo, adj.double(), p, "")
    (ref * R.double()).sum().backward()

    def rel(a, b):
        return float((a.double() - b).norm() / b.norm().clamp_min(1e-30))
    errs = {"out": rel(out.detach(), ref.detach()), "dX": rel(xr.grad, xo.grad)}
    errs.update({k: rel(v.grad, p[k].grad) for k, v in mod.named_parameters()})
    assert all(e < 2e-3 for e in errs.values()), errs
    assert errs["out"] > 2e-6, "the TF32 path was not taken"   # fp32 path: ~1e-7


def test_adjacency_view_follows_the_tensor_not_its_address():
    """ADVICE (round 1): a model that is freed and rebuilt with ANOTHER adjacency of the same size, and an adjacency edited
    in place, must not be served the old graph's CSR / CSC (ops.sparse_adjacency caches per tensor object + version)."""
    import gc
    import torch
    from msgat_b200 import models, synthetic
    from oracle import msgat_oracle as O
    torch.manual_seed(3)
    N = 31
    mod = models.GACN(3, 8, 12)
    for q in mod.parameters():   # the bare layers leave their parameters uninitialised, as the reference's do (msgat.py:206 is MSGAT's)
        torch.nn.init.normal_(q, std=0.5) if q.dim() < 2 else torch.nn.init.xavier_normal_(q)
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 3, N, 12, generator=g)
    p = {k: v.detach().double() for k, v in mod.state_dict().items()}

    def check(adj):
        out = mod(x, adj)
        ref = O.gacn(x.double(), adj.double(), p, "")
        assert float((out.double() - ref).norm() / ref.norm()) < 1e-5

    adj1 = synthetic.geometric_graph(N, mean_degree=4.0, seed=1)
    check(adj1)
    del adj1
    gc.collect()
    adj2 = synthetic.geometric_graph(N, mean_degree=4.0, seed=2)   # may land on the freed tensor's storage
    check(adj2)
    with torch.no_grad():                                          # in-place edit: same object, new version
        adj2[0, N - 1] = adj2[N - 1, 0] = 0.37
    check(adj2)
