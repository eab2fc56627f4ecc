"""Developer helper for compute-sanitizer --tool initcheck: one small grouped forward + backward."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from msgat_b200 import models, synthetic
from msgat_b200.loss import huber_loss
torch.manual_seed(0)
N, B, M = int(os.environ.get("DN", 45)), int(os.environ.get("DB", 3)), 2
adj = synthetic.road_graph(N, int(N * 1.2), seed=0)
model = models.msgat72(n_components=M, in_channels=1, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone()).cuda()
b = synthetic.make_batch(B, M, 1, N, seed=3)
out = model(b.X.cuda(), b.H.cuda(), b.D.cuda())
huber_loss(out, b.Y.cuda(), 50.0).backward()
torch.cuda.synchronize()
print("done", float(out.abs().sum()))
