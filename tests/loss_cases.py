"""Loss / metrics / optimizer kernels against the oracle's statement of loss.py:51-52, metrics.py:20-35 and
torch.optim.Adam (engine.py:106); shared by the emulated (CPU) and the GPU tests."""
import torch

from oracle import msgat_oracle as O


def run_loss_case(device, n_shape=(3, 37, 12), delta=50.0, seed=0):
    from msgat_b200 import loss as L
    g = torch.Generator().manual_seed(seed)
    true = (200.0 + 100.0 * torch.randn(*n_shape, generator=g)).clamp_min(0.0)   # zeros exercise the MAPE mask
    pred = true + 60.0 * torch.randn(*n_shape, generator=g)                      # both Huber branches fire
    true[0, 0, :3] = 0.0
    errs = {}
    p64 = pred.double().requires_grad_(True)
    ref = O.huber(p64, true.double(), delta)
    (ref * 3.0).backward()
    frac_lin = float(((pred - true).abs() > delta).float().mean())
    assert 0.05 < frac_lin < 0.95
    m = L.Metrics()
    pd = pred.to(device).requires_grad_(True)
    td = true.to(device)
    out = L.HuberLoss(delta)(pd, td, metrics=m)
    (out * 3.0).backward()
    errs["loss"] = abs(float(out) - float(ref)) / abs(float(ref))
    errs["dpred"] = float((pd.grad.cpu().double() - p64.grad).norm() / p64.grad.norm())
    # a second batch through Metrics.update alone: the sums are streaming
    pred2 = pred * 0.9 + 5.0
    m.update(pred2.to(device), td)
    st = O.metrics_update({}, pred.double(), true.double())
    st = O.metrics_update(st, pred2.double(), true.double())
    got = m.todict()
    for k in ("MAE", "MAPE", "RMSE"):
        errs[k] = abs(got[k] - st[k]) / abs(st[k])
    errs["n"] = abs(m.n - st["n"])
    # odd length: the scalar tail
    a, b = torch.randn(1031, generator=g) * 3, torch.randn(1031, generator=g)
    o2 = L.huber_loss(a.to(device), b.to(device), 1.0)
    errs["tail"] = abs(float(o2) - float(O.huber(a.double(), b.double(), 1.0))) / float(O.huber(a.double(), b.double(), 1.0))
    return errs


def run_adam_case(device, steps=4, world=2):
    from msgat_b200.distributed import FlatGradients
    from msgat_b200.optim import FlatAdam
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Linear(5, 3)).to(device)
    ref = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Linear(5, 3)).double()
    ref.load_state_dict({k: v.detach().cpu().double() for k, v in model.state_dict().items()})
    fg = FlatGradients(model)
    opt = FlatAdam(fg, lr=1e-3, weight_decay=5e-4, grad_scale=1.0 / world)
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-3, weight_decay=5e-4)
    for s in range(steps):
        x = torch.randn(11, 7)
        fg.zero()
        (model(x.to(device)) ** 2).sum().mul(world).backward()   # as if `world` ranks had been summed
        opt.step()
        ropt.zero_grad()
        (ref(x.double()) ** 2).sum().backward()
        ropt.step()
    err = 0.0
    for (n, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
        err = max(err, float((a.detach().cpu().double() - b.detach()).norm() / b.detach().norm()))
    assert int(opt.step_count.cpu()) == steps
    return err
