"""GPU: the reference's own train loop (``engine.py:40-84``, ``Trainer`` ``:101-108``, unmodified, from the git-ignored
``oracle/_ref`` copy) drives the drop-in model: autocast + GradScaler + torch Adam + the reference's HuberLoss / Metrics,
three steps, against the same three steps of the fp64 oracle model on the CPU."""
import os
import sys
import tempfile

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))

pytestmark = pytest.mark.gpu


def test_reference_engine_drives_the_dropin():
    import make_ref
    engine = make_ref.load_engine()
    if engine is None:
        pytest.skip("oracle/_ref/engine.py is not present (made by oracle/make_ref.py where /root/reference exists)")
    from msgat_b200 import models, synthetic
    from oracle import msgat_oracle as O
    torch.manual_seed(5)
    shape = synthetic.PEMS_SHAPES["pemsd8"]
    adj = synthetic.make_adjacency("pemsd8")
    model = models.msgat72(n_components=shape["M"], in_channels=shape["C"], in_timesteps=12, out_timesteps=12, use_te=True,
                           adj=adj.clone())
    p64 = {k: v.detach().clone().double() for k, v in model.state_dict().items()}
    batches = [synthetic.make_batch(4, shape["M"], shape["C"], shape["N"], seed=100 + i) for i in range(3)]
    data = [(b.X, b.H, b.D, b.Y) for b in batches]

    # the fp64 oracle under the same protocol (engine.py:106: Adam lr 1e-3, weight_decay 5e-4; loss delta 50)
    train = {k: v.requires_grad_(True) for k, v in p64.items() if k != "adj"}
    opt = torch.optim.Adam(list(train.values()), lr=1e-3, weight_decay=5e-4)
    ref_losses = []
    for X, H, D, Y in data:
        pred = O.msgat(X.double(), H, D, p64, O.MODEL_DILATIONS["msgat72"])
        loss = O.huber(pred, Y.double(), 50.0)
        opt.zero_grad()
        loss.backward()
        opt.step()
        ref_losses.append(float(loss))

    with tempfile.TemporaryDirectory() as out_dir:
        trainer = engine.Trainer(model.cuda(), loss_delta=50.0, out_dir=out_dir)
        loss_ave = trainer.run_epoch(data, gpu_id=0, epoch=0, mode="train")     # engine.py:40
        log = open(os.path.join(out_dir, "run.log")).read()
    assert "[Train   ]" in log and "MAE=" in log
    assert abs(loss_ave - sum(ref_losses) / 3) / (sum(ref_losses) / 3) < 1e-4, (loss_ave, ref_losses)
    # parameters after three optimizer steps taken by the reference's GradScaler / Adam on the drop-in's gradients
    worst = 0.0
    for k, v in model.state_dict().items():
        if k == "adj":
            continue
        d = (v.detach().cpu().double() - p64[k].detach()).norm() / p64[k].detach().norm().clamp_min(1e-30)
        worst = max(worst, float(d))
    # calibration: the fp32 oracle against the fp64 oracle under the same protocol on the CPU differs by 1.5e-4 on its
    # worst tensor (Adam's first steps are lr * sign-like: near-zero gradient coordinates amplify rounding noise)
    assert worst < 1e-3, worst
