"""Parity of the sm_100a kernels (through the drop-in modules -> ctypes -> C ABI) with the
reference: golden fixtures minted from the unmodified reference, the CPU oracle on seeded
PEMS-shaped inputs, and size-independent properties at the full benchmark shapes."""
import pytest
import torch

from conftest import Golden, golden_cases, rel_l2
from layer_cases import run_case
from parity import (FP32_TOL, check_case, is_attention_scalar, record, relu_conditioning, tensor_class,
                    tolerance)

pytestmark = pytest.mark.gpu

from msgat_b200.loss import huber_loss  # noqa: E402  (the product's device-side loss; the oracle keeps O.huber)


@pytest.mark.parametrize("case", golden_cases(""))
def test_kernels_match_reference_golden(case):
    G = Golden(case)
    out, grads, gin = run_case(G, device="cuda")
    fails, _ = check_case(G, out, grads, gin)
    if fails:  # is the reference itself that sensitive to sign flips of near-zero ReLU inputs?
        _note(f"[relu-conditioning probe] {case}: " + "; ".join(fails))
        fails, _ = check_case(G, out, grads, gin, cond=_golden_conditioning(G))
    assert not fails, "\n".join(fails)


def _note(msg):
    """Cases that needed the ReLU-conditioning probe are listed in gpurun_out/parity_notes.log (when that directory exists)."""
    import os
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_notes.log"), "a") as f:
            f.write(msg + "\n")
    print(msg)


def _golden_conditioning(G):
    from oracle_calls import oracle_forward

    def run():
        p = {k: v.double().requires_grad_(v.is_floating_point() and k != "adj") for k, v in G.params.items()}
        ins = {k: (v.double().requires_grad_(True) if v.is_floating_point() else v) for k, v in G.inputs.items()}
        out = oracle_forward(G.case, G.meta, p, ins)
        (out * G.R.double()).sum().backward()
        res = {"out": out.detach()}
        res.update({"grad " + n: p[n].grad for n in G.grads})
        res.update({"dinput " + n: ins[n].grad for n in G.gin if n != "adjacency"})
        return res
    return relu_conditioning(run)


def compare_with_oracle(model, out, X, ref, ref32, X64, X32, p, p32, floor=FP32_TOL, cond=None):
    """rel-L2 of out, dX and every parameter gradient against the fp64 oracle under the
    tolerance protocol of tests/parity.py (noise = the oracle's own fp32 run)."""
    rows = [("out", out, ref, ref32), ("dX", X.grad, X64.grad, X32.grad)]
    for n, prm in model.named_parameters():
        if p[n].grad is not None:
            rows.append((n, prm.grad, p[n].grad, p32[n].grad))
    noises = [rel_l2(w32, w) for _, _, w, w32 in rows]
    envelope = max(noises)
    fails, pending = [], []
    for (label, got, want, _), n in zip(rows, noises):
        scalar = is_attention_scalar(label)
        e, tol = rel_l2(got, want), tolerance(n, envelope, floor, scalar=scalar)
        if cond is not None:
            tol = max(tol, 2.0 * cond.get(label, 0.0))
        pending.append((tensor_class(label if label in ("out", "dX") else "grad " + label) + (" [probe]" if cond is not None else ""), e, n, tol))
        if not e < tol:
            fails.append(f"{label}: {e:.3e} > {tol:.2e} (fp32 oracle noise {n:.2e}, case envelope {envelope:.2e})")
    if not fails:   # margins of a pass that succeeded; a failing first pass is re-judged with the probe (tests/parity.py)
        for r in pending:
            record(*r)
    return fails


def _model_and_batch(name, B, width="msgat72", seed=0, N=None):
    from msgat_b200 import models, synthetic
    s = dict(synthetic.PEMS_SHAPES[name])
    if N is not None:
        s["N"], s["E"] = N, int(N * 1.2)
    torch.manual_seed(seed)
    adj = synthetic.road_graph(s["N"], s["E"], seed=0) if s["E"] else synthetic.geometric_graph(s["N"])
    model = getattr(models, width)(n_components=s["M"], in_channels=s["C"], in_timesteps=12, out_timesteps=12,
                                   use_te=True, adj=adj.clone())
    batch = synthetic.make_batch(B, s["M"], s["C"], s["N"], seed=1234)
    return model, batch


@pytest.mark.parametrize("name,B", [("pemsd8", 3), ("pemsd4", 2), ("pemsd3", 2), ("pemsd7", 2)])
def test_full_model_matches_oracle_fp64(name, B):
    """Forward, dX and every parameter gradient of the whole MSGAT at the real graph sizes
    against the CPU oracle in fp64 (same state_dict)."""
    from oracle import msgat_oracle as O
    model, batch = _model_and_batch(name, B)
    p = {k: v.detach().double().requires_grad_(v.is_floating_point() and k != "adj")
         for k, v in model.state_dict().items()}
    X64 = batch.X.double().requires_grad_(True)
    ref = O.msgat(X64, batch.H, batch.D, p, O.MODEL_DILATIONS["msgat72"])
    O.huber(ref, batch.Y.double(), 50.0).backward()
    # the reference arithmetic in fp32 on the same inputs gives the noise floor of each tensor
    p32 = {k: v.detach().float().requires_grad_(v.is_floating_point() and k != "adj") for k, v in p.items()}
    X32 = batch.X.clone().requires_grad_(True)
    ref32 = O.msgat(X32, batch.H, batch.D, p32, O.MODEL_DILATIONS["msgat72"])
    O.huber(ref32, batch.Y, 50.0).backward()

    model = model.cuda()
    X = batch.X.cuda().requires_grad_(True)
    out = model(X, batch.H.cuda(), batch.D.cuda())
    huber_loss(out, batch.Y.cuda(), 50.0).backward()
    fails = compare_with_oracle(model, out, X, ref, ref32, X64, X32, p, p32)
    if fails:  # is the reference itself that sensitive to sign flips of near-zero ReLU inputs?
        _note(f"[relu-conditioning probe] full model {name}: " + "; ".join(fails))

        def run():
            q = {k: v.detach().clone().requires_grad_(v.requires_grad) for k, v in p.items()}
            Xq = batch.X.double().requires_grad_(True)
            r = O.msgat(Xq, batch.H, batch.D, q, O.MODEL_DILATIONS["msgat72"])
            O.huber(r, batch.Y.double(), 50.0).backward()
            res = {"out": r.detach(), "dX": Xq.grad}
            res.update({n: q[n].grad for n in q if q[n].grad is not None})
            return res
        fails = compare_with_oracle(model, out, X, ref, ref32, X64, X32, p, p32, cond=relu_conditioning(run))
    assert not fails, "\n".join(fails)


def test_pemsd7_full_batch_properties():
    """At BASELINE.json's full size (N=883, B=32) the oracle is too slow; check properties:
    every sample of the batch is computed independently (bit-exact vs a batch of one),
    the result is deterministic, and the loss gradient is finite everywhere."""
    from oracle import msgat_oracle as O
    model, batch = _model_and_batch("pemsd7", 32)
    model = model.cuda()
    X, H, D, Y = batch.X.cuda(), batch.H.cuda(), batch.D.cuda(), batch.Y.cuda()
    Xg = X.clone().requires_grad_(True)
    out = model(Xg, H, D)
    huber_loss(out, Y, 50.0).backward()
    assert torch.isfinite(out).all() and torch.isfinite(Xg.grad).all()
    for prm in model.parameters():
        if prm.grad is not None:
            assert torch.isfinite(prm.grad).all()
    with torch.no_grad():
        out2 = model(X, H, D)
        assert torch.equal(out2, out.detach()), "forward is not deterministic"
        for b in (0, 17, 31):
            ob = model(X[b:b + 1], H[b:b + 1], D[b:b + 1])
            assert torch.equal(ob[0], out2[b]), f"sample {b} depends on its batch"


def test_node_permutation_equivariance():
    """Relabelling the sensors (adjacency, per-node parameters, inputs) permutes the output:
    attention over N is permutation-equivariant (attention.py:32-36)."""
    from msgat_b200 import models
    model, batch = _model_and_batch("pemsd8", 2)
    N = model.adj.shape[0]
    perm = torch.randperm(N, generator=torch.Generator().manual_seed(5))
    sd = model.state_dict()
    sd2 = {}
    M, Tout = len(model.tpcs), 12
    for k, v in sd.items():
        if k == "adj":
            sd2[k] = v[perm][:, perm]
        elif k.endswith(("cacn.seq.0.alpha",)):
            sd2[k] = v[perm]
        elif k.endswith(("Wt1", "Wt2")):
            sd2[k] = v[:, perm]
        elif k.startswith("te."):
            sd2[k] = v.view(v.shape[0], M, N, Tout)[:, :, perm].reshape(v.shape)
        else:
            sd2[k] = v
    model2 = models.msgat72(n_components=M, in_channels=batch.X.shape[2], in_timesteps=12, out_timesteps=12,
                            use_te=True, adj=sd2["adj"].clone())
    model2.load_state_dict(sd2)
    model, model2 = model.cuda(), model2.cuda()
    with torch.no_grad():
        a = model(batch.X.cuda(), batch.H.cuda(), batch.D.cuda())
        b = model2(batch.X[:, :, :, perm].cuda(), batch.H.cuda(), batch.D.cuda())
    assert rel_l2(b, a[:, perm]) < 1e-5


def test_engine_protocol_autocast_and_state_dict():
    """The reference engine calls the model under cuda.amp.autocast with a GradScaler
    (engine.py:54,61) and round-trips state_dict (engine.py:139,152)."""
    from oracle import msgat_oracle as O
    model, batch = _model_and_batch("pemsd8", 2)
    model = model.cuda()
    scaler = torch.amp.GradScaler("cuda")
    with torch.autocast("cuda"):
        out = model(batch.X.cuda(), batch.H.cuda(), batch.D.cuda())
        loss = huber_loss(out.float(), batch.Y.cuda(), 50.0)
    scaler.scale(loss).backward()
    assert out.dtype == torch.float32
    with torch.no_grad():
        ref = model(batch.X.cuda(), batch.H.cuda(), batch.D.cuda())
    assert torch.equal(ref, out.detach()), "autocast changed the arithmetic of the fp32 path"
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    model.load_state_dict(sd)


@pytest.mark.gpu
def test_n4096_full_size_properties():
    """BASELINE.json's largest configuration (N=4096, C=3) at full graph size: the oracle's dense
    [B,N,N] attention is too slow to check here, so the test uses properties that do not depend on the
    size: finite outputs and gradients, bit-reproducibility, and independence of a sample from the rest
    of its batch (every operator of the path is per-sample)."""
    from msgat_b200 import models, synthetic
    shape = synthetic.PEMS_SHAPES["n4096"]
    torch.manual_seed(3)
    adj = synthetic.make_adjacency("n4096")
    model = models.msgat72(n_components=2, in_channels=shape["C"], in_timesteps=12, out_timesteps=12, use_te=True,
                           adj=adj.clone()).cuda()
    b = synthetic.make_batch(2, 2, shape["C"], shape["N"], seed=11)
    X, H, D, Y = (t.cuda() for t in (b.X, b.H, b.D, b.Y))

    def run(Xs, Hs, Ds, Ys):
        model.zero_grad(set_to_none=True)
        Xs = Xs.clone().requires_grad_(True)
        out = model(Xs, Hs, Ds)
        huber_loss(out, Ys, 50.0).backward()
        torch.cuda.synchronize()
        return out.detach(), Xs.grad, [p.grad.clone() for p in model.parameters() if p.grad is not None]

    o1, gx1, gp1 = run(X, H, D, Y)
    o2, gx2, gp2 = run(X, H, D, Y)
    assert torch.isfinite(o1).all() and torch.isfinite(gx1).all() and all(torch.isfinite(g).all() for g in gp1)
    assert torch.equal(o1, o2) and torch.equal(gx1, gx2) and all(torch.equal(a, c) for a, c in zip(gp1, gp2))
    o_single, _, _ = run(X[1:], H[1:], D[1:], Y[1:])
    assert torch.equal(o_single[0], o1[1]), "a sample's output depends on the rest of its batch"
    assert float(gx1.abs().max()) > 0


def test_static_gate_use_te_false_matches_oracle():
    """use_te=False: the reference registers a static gate W [M,N,Tout] (msgat.py:186-189) but its forward
    dereferences self.te and crashes; the drop-in applies W as the gate,
    out = sum_m tpc_m(X[:, m]) * W[m].  Checked against the oracle's TPC with that gate, incl. dW."""
    from msgat_b200 import models, synthetic
    from oracle import msgat_oracle as O
    torch.manual_seed(4)
    N, C, M, B = 33, 3, 2, 2
    adj = synthetic.road_graph(N, 40, seed=2)
    model = models.msgat72(n_components=M, in_channels=C, in_timesteps=12, out_timesteps=12, use_te=False, adj=adj.clone())
    batch = synthetic.make_batch(B, M, C, N, seed=5)
    res = {}
    for dt in (torch.float64, torch.float32):
        p = {k: v.detach().clone().to(dt).requires_grad_(v.is_floating_point() and k != "adj") for k, v in model.state_dict().items()}
        X = batch.X.detach().clone().to(dt).requires_grad_(True)
        ref = sum(O.tpc(X[:, m], p["adj"], p, f"tpcs.{m}.", O.MODEL_DILATIONS["msgat72"]) * p["W"][m] for m in range(M))
        O.huber(ref, batch.Y.to(dt), 50.0).backward()
        res[dt] = (ref.detach(), X.grad, p)
    model = model.cuda()
    Xg = batch.X.cuda().requires_grad_(True)
    out = model(Xg, batch.H.cuda(), batch.D.cuda())
    huber_loss(out, batch.Y.cuda(), 50.0).backward()
    ref, dX, p = res[torch.float64]
    ref32, dX32, p32 = res[torch.float32]
    fails = compare_with_oracle(model, out, Xg, ref, ref32, type("T", (), {"grad": dX}), type("T", (), {"grad": dX32}), p, p32)
    assert not fails, "\n".join(fails)
    assert model.W.grad is not None and float(model.W.grad.abs().max()) > 0


def _ref_init(module, seed):
    """MSGAT.reset_parameters rule (msgat.py:206-217) on a sub-module."""
    torch.manual_seed(seed)
    for prm in module.parameters():
        if prm.ndim >= 2:
            torch.nn.init.xavier_normal_(prm)
        else:
            s = prm.size(0) ** -0.5
            torch.nn.init.uniform_(prm, -s, s)


def _meam_vs_oracle(C, N, adj, B, dil, precision, floor, seed=0, res_bias_shift=0.0, peaked=False):
    """One MEAM (msgat.py:121-131) on the device against the oracle in fp64 (+ fp32 for the noise):
    -> list of failures under the per-tensor protocol with the given floor."""
    from msgat_b200 import models
    from oracle import msgat_oracle as O
    mod = models.MEAM(C, 72, N, 12, dil)
    _ref_init(mod, seed)
    with torch.no_grad():
        mod.res.bias.add_(res_bias_shift)
        if peaked:
            mod.gacn.gatt.Wg.mul_(30.0); mod.cacn.seq[0].Wc.mul_(30.0)
            mod.tacn.seq[0].Wt1.mul_(5.0); mod.tacn.seq[0].Wt2.mul_(5.0)
    mod.precision = precision
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(B, C, N, 12, generator=g)
    R = torch.randn(B, 72, N, 12, generator=g)
    res = {}
    for dt in (torch.float64, torch.float32):
        p = {k: v.detach().clone().to(dt).requires_grad_(True) for k, v in mod.state_dict().items()}
        xr = x.detach().clone().to(dt).requires_grad_(True)
        ref = O.meam(xr, adj.to(dt), p, "", dil)
        (ref * R.to(dt)).sum().backward()
        res[dt] = (ref.detach(), xr.grad, p)
    mod = mod.cuda()
    xg = x.cuda().requires_grad_(True)
    out = mod(xg, adj.cuda())
    (out * R.cuda()).sum().backward()
    ref, dX, p = res[torch.float64]
    ref32, dX32, p32 = res[torch.float32]
    fails = compare_with_oracle(mod, out, xg, ref, ref32, type("T", (), {"grad": dX}), type("T", (), {"grad": dX32}),
                                p, p32, floor=floor)
    return fails, float(ref.min())


@pytest.mark.parametrize("C", [3, 72])
def test_n4096_meam_matches_oracle_fp64(C):
    """BASELINE.json config 5 at FULL graph size against the oracle (attention.py:32-36 with a dense
    [1,4096,4096] fp64 attention): one MEAM, B=1, the distance-thresholded graph, C=3 (layer 1) and C=72 (layer 2).
    k_sstats / k_scol run 16 column chunks here and the CSR rows are ~9 wide."""
    from msgat_b200 import _lib, synthetic
    adj = synthetic.geometric_graph(4096)
    fails, _ = _meam_vs_oracle(C, 4096, adj, 1, [2, 4] if C == 72 else [1, 2], _lib.PREC_FP32, FP32_TOL, seed=C)
    assert not fails, "\n".join(fails)


def test_n4096_full_model_matches_oracle_fp64():
    """msgat72 with one component, B=1, N=4096: out, dX and every parameter gradient against the fp64 oracle."""
    from msgat_b200 import models, synthetic
    from oracle import msgat_oracle as O
    torch.manual_seed(9)
    adj = synthetic.geometric_graph(4096)
    model = models.msgat72(n_components=1, in_channels=3, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone())
    batch = synthetic.make_batch(1, 1, 3, 4096, seed=21)
    res = {}
    for dt in (torch.float64, torch.float32):
        p = {k: v.detach().clone().to(dt).requires_grad_(v.is_floating_point() and k != "adj") for k, v in model.state_dict().items()}
        X = batch.X.detach().clone().to(dt).requires_grad_(True)
        ref = O.msgat(X, batch.H, batch.D, p, O.MODEL_DILATIONS["msgat72"])
        O.huber(ref, batch.Y.to(dt), 50.0).backward()
        res[dt] = (ref.detach(), X, p)
    model = model.cuda()
    Xg = batch.X.cuda().requires_grad_(True)
    out = model(Xg, batch.H.cuda(), batch.D.cuda())
    huber_loss(out, batch.Y.cuda(), 50.0).backward()
    fails = compare_with_oracle(model, out, Xg, res[torch.float64][0], res[torch.float32][0], res[torch.float64][1],
                                res[torch.float32][1], res[torch.float64][2], res[torch.float32][2])
    assert not fails, "\n".join(fails)
