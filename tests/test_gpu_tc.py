"""tcgen05 (kind::tf32) building block of the tensor-core path: the SW128 tile layout, the
shared-memory / instruction descriptors and the TMEM read-back, against torch fp64."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tf32(x):
    """round-to-nearest-even drop of the 13 low mantissa bits is an upper bound of what the
    tensor core does (it truncates); used only to size the tolerance."""
    return x


@pytest.mark.parametrize("mode", [2, 3])
@pytest.mark.parametrize("K,N", [(72, 96), (24, 32), (48, 32), (8, 16)])
def test_mn_major_channel_mix(K, N, mode):
    from msgat_b200 import _lib
    lib = _lib.get()
    g = torch.Generator().manual_seed(K * 100 + N)
    A = torch.randn(K, 256, generator=g).cuda()
    B = torch.randn(K, N, generator=g).cuda()
    D = torch.zeros(256, N, device="cuda")
    _lib.check(lib.msgat_tc_probe(mode, _lib.ptr(A), _lib.ptr(B), _lib.ptr(D), K, N, 0, 0, 0, _lib.stream_of(A)), "tc_probe")
    torch.cuda.synchronize()
    ref = A.double().T @ B.double()
    err = float((D.double() - ref).norm() / ref.norm())
    assert err < 2e-3, err


@pytest.mark.parametrize("mode", [1, 4])
@pytest.mark.parametrize("Ra,Rb,P,N", [(72, 72, 192, 80), (96, 72, 192, 80), (24, 48, 64, 48), (72, 72, 32, 80)])
def test_k_major_weight_gradient(Ra, Rb, P, N, mode):
    from msgat_b200 import _lib
    lib = _lib.get()
    g = torch.Generator().manual_seed(Ra + Rb + P)
    A = torch.randn(Ra, P, generator=g).cuda()
    B = torch.randn(Rb, P, generator=g).cuda()
    D = torch.zeros(128, N, device="cuda")
    _lib.check(lib.msgat_tc_probe(mode, _lib.ptr(A), _lib.ptr(B), _lib.ptr(D), 0, N, Ra, Rb, P, _lib.stream_of(A)), "tc_probe")
    torch.cuda.synchronize()
    ref = A.double() @ B.double().T
    got = D[:Ra, :min(Rb, N)].double()
    ref = ref[:, :min(Rb, N)]
    err = float((got - ref).norm() / ref.norm())
    assert err < 2e-3, err


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def test_tc_meam_layer2_matches_reference_golden():
    """MEAM at the layer-2 shape (C = 72) on the tcgen05 TF32 path against the golden vectors
    minted from the reference: 2e-3 relative (BASELINE.json north_star)."""
    from conftest import Golden
    from layer_cases import build_module
    from msgat_b200 import _lib
    from parity import TC_TOL
    G = Golden("meam_N29C72_peaked_d24")
    mod = build_module(G).cuda()
    mod.precision = _lib.PREC_TF32
    x = G.inputs["signals"].float().cuda().requires_grad_(True)
    adj = G.inputs["adjacency"].float().cuda()
    l0 = _lib.launch_count()
    out = mod(x, adj)
    (out * G.R.float().cuda()).sum().backward()
    assert _lib.launch_count() > l0
    assert _rel(out, G.out) < TC_TOL, _rel(out, G.out)
    print("out", _rel(out, G.out), "dx", _rel(x.grad, G.gin["signals"]))
    print(sorted(((_rel(p.grad, G.grads[n]), n) for n, p in mod.named_parameters() if n in G.grads), reverse=True)[:6])
    worst = max((_rel(p.grad, G.grads[n]), n) for n, p in mod.named_parameters() if n in G.grads)
    assert worst[0] < 10 * TC_TOL, worst


@pytest.mark.parametrize("name,B", [("pemsd8", 3), ("pemsd7", 2)])
def test_tc_full_model_matches_oracle(name, B):
    from msgat_b200 import _lib
    from oracle import msgat_oracle as O
    from parity import TC_TOL
    from test_gpu_parity import _model_and_batch
    model, batch = _model_and_batch(name, B)
    p = {k: v.detach().double().requires_grad_(v.is_floating_point() and k != "adj")
         for k, v in model.state_dict().items()}
    X64 = batch.X.double().requires_grad_(True)
    ref = O.msgat(X64, batch.H, batch.D, p, O.MODEL_DILATIONS["msgat72"])
    O.huber(ref, batch.Y.double(), 50.0).backward()
    model = model.cuda().set_precision(_lib.PREC_TF32)
    X = batch.X.cuda().requires_grad_(True)
    out = model(X, batch.H.cuda(), batch.D.cuda())
    O.huber(out, batch.Y.cuda(), 50.0).backward()
    assert _rel(out, ref) < TC_TOL, _rel(out, ref)
    print("out", _rel(out, ref), "dX", _rel(X.grad, X64.grad))
    rows = sorted(((_rel(prm.grad, p[n].grad), n) for n, prm in model.named_parameters() if p[n].grad is not None
                   and float(p[n].grad.norm()) > 0), reverse=True)
    print("worst TF32-path gradients:", rows[:5])
    assert rows[0][0] < 10 * TC_TOL, rows[:5]
