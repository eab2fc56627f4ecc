"""Developer helper: repeat the forward of one grouped layer-2 MEAM and count bitwise mismatches."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from msgat_b200 import models, synthetic
torch.manual_seed(0)
N = int(os.environ.get("DN", 883))
adj = synthetic.road_graph(N, int(N * 1.2), seed=0)
for B, M in [(1, 1), (1, 5), (4, 5), (32, 5)]:
    model = models.msgat72(n_components=M, in_channels=1, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone()).cuda()
    b = synthetic.make_batch(B, M, 1, N, seed=3)
    X, H, D = b.X.cuda(), b.H.cuda(), b.D.cuda()
    with torch.no_grad():
        ref = model(X, H, D).clone()
        bad = 0
        for i in range(6):
            # poison the caching allocator's free blocks: a read of uninitialised memory now sees NaNs / other values
            junk = torch.full((500_000_000,), float("nan") if i % 2 else 1e30, device="cuda"); del junk
            o = model(X, H, D)
            if not torch.equal(o, ref):
                bad += 1
                d = (o - ref).abs()
                idx = d.flatten().argmax()
                print(f"   mismatch B={B} M={M} run {i}: max|diff| {float(d.max()):.3e} n_diff {int((d > 0).sum())} at flat {int(idx)}")
    print(f"B={B} M={M}: {bad}/6 mismatching forwards")
