"""Developer helper: the scenario of test_pemsd7_full_batch_properties (forward + backward, then forward again)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from msgat_b200 import models, synthetic
from msgat_b200.loss import huber_loss
poison = os.environ.get("POISON", "0") == "1"
def scribble(val):
    bufs = [torch.full((n,), val, device="cuda") for n in (300_000_000, 60_000_000, 60_000_000, 25_000_000, 25_000_000, 10_000_000, 3_000_000, 1_000_000, 300_000, 40_000) for _ in range(2)]
    del bufs
adj = synthetic.road_graph(883, 866, seed=0)
fails = 0
for it in range(6):
    torch.manual_seed(it)
    model = models.msgat72(n_components=5, in_channels=1, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone()).cuda()
    b = synthetic.make_batch(32, 5, 1, 883, seed=1234)
    X, H, D, Y = b.X.cuda(), b.H.cuda(), b.D.cuda(), b.Y.cuda()
    if poison: scribble(float("nan"))
    Xg = X.clone().requires_grad_(True)
    out = model(Xg, H, D)
    huber_loss(out, Y, 50.0).backward()
    if poison: scribble(1e30)
    with torch.no_grad():
        out2 = model(X, H, D)
    torch.cuda.synchronize()
    if not torch.equal(out2, out.detach()):
        fails += 1
        d = (out2 - out.detach()).abs()
        nz = (d > 0).nonzero()
        print(f"  iter {it}: n_diff {len(nz)} max {float(d.max()):.3e} nan1 {int(torch.isnan(out).sum())} nan2 {int(torch.isnan(out2).sum())} first idx {nz[0].tolist()} last {nz[-1].tolist()}")
    del model
print(f"poison={poison}: {fails}/6 failed")
