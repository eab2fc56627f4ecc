"""GPU: the CUDA-graph replay of a training step is the eager step, launch for launch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(seed, direct=True):
    from msgat_b200 import models, synthetic
    from msgat_b200.distributed import FlatGradients
    from msgat_b200.optim import FlatAdam
    shape = synthetic.PEMS_SHAPES["pemsd8"]
    torch.manual_seed(seed)
    adj = synthetic.make_adjacency("pemsd8")
    model = models.msgat72(n_components=shape["M"], in_channels=shape["C"], in_timesteps=12, out_timesteps=12, use_te=True,
                           adj=adj.clone()).cuda()
    fg = FlatGradients(model, direct=direct)
    opt = FlatAdam(fg, lr=1e-3, weight_decay=5e-4)
    b = synthetic.make_batch(3, shape["M"], shape["C"], shape["N"], seed=7)
    return model, fg, opt, [t.cuda() for t in (b.X, b.H, b.D, b.Y)]


def test_graphed_step_matches_eager_steps():
    from msgat_b200.train import GraphedStep
    from msgat_b200.loss import huber_loss
    loss_fn = lambda o, y: huber_loss(o, y, 50.0)  # noqa: E731
    m1, fg1, opt1, batch = _setup(0)
    m2, fg2, opt2, _ = _setup(0)
    g = GraphedStep(m1, loss_fn, opt1, fg1, batch, warmup=2)      # 2 warm-up steps + 1 captured (not run)
    e = GraphedStep.__new__(GraphedStep)
    e.model, e.loss_fn, e.opt, e.grads = m2, loss_fn, opt2, fg2
    for _ in range(2):
        e.eager(*batch)
    for _ in range(3):
        lg = float(g(*batch))
        le = float(e.eager(*batch))
        assert lg == le
    torch.cuda.synchronize()
    for (n, a), (_, b) in zip(m1.named_parameters(), m2.named_parameters()):
        assert torch.equal(a, b), n


def test_gradient_sinks_equal_autograd_accumulation():
    """FlatGradients(direct=True): the kernels write parameter gradients straight into the flat buffer; the result is
    bit-identical to letting autograd accumulate returned gradients into the zeroed buffer."""
    from msgat_b200.loss import huber_loss
    outs = []
    for direct in (True, False):
        model, fg, _, batch = _setup(3, direct=direct)
        fg.zero()
        huber_loss(model(*batch[:3]), batch[3], 50.0).backward()
        torch.cuda.synchronize()
        outs.append(fg.flat.clone())
    assert torch.equal(outs[0], outs[1])


_FORK_PROBE = r"""
import hashlib, sys, torch
sys.path.insert(0, {root!r})
from msgat_b200 import models, synthetic
from msgat_b200.loss import huber_loss
torch.manual_seed(5)
N, C, M, B = 170, 3, 3, 4
adj = synthetic.road_graph(N, 200, seed=2)
model = models.msgat72(n_components=M, in_channels=C, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone()).cuda()
b = synthetic.make_batch(B, M, C, N, seed=11)
h = hashlib.sha256()
for _ in range(3):
    model.zero_grad(set_to_none=True)
    out = model(b.X.cuda(), b.H.cuda(), b.D.cuda())
    huber_loss(out, b.Y.cuda(), 50.0).backward()
    torch.cuda.synchronize()
    h.update(out.detach().cpu().numpy().tobytes())
    for p in model.parameters():
        if p.grad is not None:
            h.update(p.grad.cpu().numpy().tobytes())
print("DIGEST", h.hexdigest())
"""


def test_side_stream_changes_no_bit():
    """The library forks a side stream for the per-sample kernels of a MEAM (api.cu: msgat_fork / msgat_join).  With
    MSGAT_NO_FORK=1 everything runs on the caller's stream: outputs and every parameter gradient of three training steps
    must be bit-identical either way (a missed dependency between the two branches would show here as a difference)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    digests = []
    for no_fork in (False, True):
        env = dict(os.environ)
        env.pop("MSGAT_NO_FORK", None)
        if no_fork:
            env["MSGAT_NO_FORK"] = "1"
        r = subprocess.run([sys.executable, "-c", _FORK_PROBE.format(root=root)], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        digests.append([l for l in r.stdout.splitlines() if l.startswith("DIGEST")][-1])
    assert digests[0] == digests[1]
