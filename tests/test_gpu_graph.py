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
