"""TPC head (msgat.py:158-160) + time gate (msgat.py:204) against a plain fp64 PyTorch statement of the
reference ops, on shapes the golden cases do not reach: node counts that are not a multiple of the 64-node
block of the streaming kernels, channel counts that are not a multiple of the 8 warps, a missing gate /
accumulator, and Tout != 12 (the generic kernels)."""
import torch

HEAD_CASES = [  # B, C, N, Tout, gate, acc, x_is_relu
    (3, 5, 70, 12, True, True, False),
    (2, 72, 130, 12, True, False, True),
    (2, 9, 33, 12, False, False, True),
    (1, 24, 64, 12, True, True, True),
    (2, 9, 33, 6, True, True, True),
]


def reference_head(x, ln_w, ln_b, fc_w, fc_b, gate, acc):
    """LayerNorm over T, Conv2d(T -> Tout, kernel [1, C]) on the transposed tensor, gate, running sum."""
    h = torch.nn.functional.layer_norm(x, (12,), ln_w, ln_b, 1e-5)
    y = torch.nn.functional.conv2d(h.transpose(1, 3), fc_w, fc_b)[..., 0].transpose(1, 2)   # [B, N, Tout]
    if gate is not None:
        y = y * gate
    return y if acc is None else acc + y


def run_head_case(case, device):
    from msgat_b200 import ops
    B, C, N, Tout, has_gate, has_acc, relu = case
    g = torch.Generator().manual_seed(1000 + 7 * C + N + Tout)
    x = torch.randn(B, C, N, 12, generator=g)
    if relu:
        x = torch.relu(x)
    t = dict(x=x, ln_w=torch.rand(12, generator=g) - 0.5, ln_b=torch.rand(12, generator=g) - 0.5,
             fc_w=torch.randn(Tout, 12, 1, C, generator=g) * 0.2, fc_b=torch.randn(Tout, generator=g),
             gate=torch.randn(B, N, Tout, generator=g) if has_gate else None,
             acc=torch.randn(B, N, Tout, generator=g) if has_acc else None)
    dout = torch.randn(B, N, Tout, generator=g)
    ref_in = {k: (None if v is None else v.double().requires_grad_(True)) for k, v in t.items()}
    ref = reference_head(*[ref_in[k] for k in ("x", "ln_w", "ln_b", "fc_w", "fc_b", "gate", "acc")])
    ref.backward(dout.double())
    dev_in = {k: (None if v is None else v.to(device).requires_grad_(True)) for k, v in t.items()}
    out = ops.head_op(dev_in["x"], dev_in["ln_w"], dev_in["ln_b"], dev_in["fc_w"], dev_in["fc_b"],
                      gate=dev_in["gate"], acc=dev_in["acc"], x_is_relu=relu)
    out.backward(dout.to(device))
    errs = {"out": rel(out, ref)}
    for k, v in dev_in.items():
        if v is None:
            continue
        gref = ref_in[k].grad
        if k == "x" and relu:   # the kernel folds the ReLU that produced x into dx (msgat.py:131)
            gref = gref * (ref_in[k] > 0)
        errs["d" + k] = rel(v.grad, gref)
    return errs


def rel(a, b):
    return float((a.detach().cpu().double() - b.detach()).norm() / b.detach().norm().clamp_min(1e-300))
