"""Developer helper: repeat the grouped layer-2 MEAM forward, localise bitwise differences (component, sample, channels, nodes)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from msgat_b200 import models, ops, synthetic
torch.manual_seed(0)
N, B, M = 883, 32, 5
adj = synthetic.road_graph(N, 866, seed=0)
model = models.msgat72(n_components=M, in_channels=1, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone()).cuda()
sp = ops.sparse_adjacency(model.adj)
x = torch.relu(torch.randn(M, B, 72, N, 12, device="cuda"))
meams = [t.tgacns[1] for t in model.tpcs]
cfg = meams[0].config()
def run():
    with torch.no_grad():
        return ops.meam_group_op(x, "mb", cfg, sp, [mm.slots() for mm in meams])
ref = run().clone()
nbad = 0
a = torch.randn(4096, 4096, device="cuda")
for i in range(40):
    if i % 3 == 1:
        for _ in range(3): a = (a @ a).clamp(-1, 1)      # perturb clocks / caches
    if i % 3 == 2:
        torch.cuda.synchronize()
    o = run()
    d = (o != ref)
    if d.any():
        nbad += 1
        idx = d.nonzero()
        ms, bs = idx[:, 0].unique().tolist(), idx[:, 1].unique().tolist()
        chans, nodes = idx[:, 2].unique().tolist(), idx[:, 3].unique().tolist()
        print(f"run {i}: {len(idx)} elems differ; comps {ms} samples {bs} nodes {nodes} channels {chans[:30]}{'...' if len(chans) > 30 else ''} max {float((o - ref).abs().max()):.3e}")
print(f"{nbad}/40 runs differ")
