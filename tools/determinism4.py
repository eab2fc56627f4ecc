"""Developer helper: like determinism3, but also compares the forward's `saved` buffer (which intermediate diverges first?)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from msgat_b200 import models, ops, synthetic
torch.manual_seed(0)
N, B, M, C, b = 883, 32, 5, 72, 24
adj = synthetic.road_graph(N, 866, seed=0)
model = models.msgat72(n_components=M, in_channels=1, in_timesteps=12, out_timesteps=12, use_te=True, adj=adj.clone()).cuda()
sp = ops.sparse_adjacency(model.adj)
x = torch.relu(torch.randn(M, B, 72, N, 12, device="cuda")).requires_grad_(True)
meams = [t.tgacns[1] for t in model.tpcs]
cfg = meams[0].config()
def rup4(v): return (v + 3) & ~3
BN = B * N
offs, o = {}, 0
for name, sz in [("qs", BN * 12), ("qt", BN * 12), ("Z", BN * 12 * b), ("m", BN), ("l", 2 * BN), ("E", 2 * BN * 12), ("U", BN * 12),
                 ("Y1", BN * 12 * b), ("A2", BN * 12 * 2 * b), ("pw", B * sp.nnz), ("qc", B * C * 12), ("Pc", B * C * C), ("Mc", B * b * C),
                 ("K", B * 240), ("Pt", B * 144)]:
    offs[name] = o; o += rup4(sz)
total = rup4(o)
def run():
    out = ops.meam_group_op(x, "mb", cfg, sp, [mm.slots() for mm in meams])
    saved = out.grad_fn.saved_tensors[2]
    return out.detach(), saved
ref, sref = run()
ref, sref = ref.clone(), sref.clone()
assert sref.numel() == total * M, (sref.numel(), total * M)
nbad = 0
for i in range(120):
    o, s = run()
    if (o != ref).any() or (s.view(torch.int32) != sref.view(torch.int32)).any():
        nbad += 1
        idx = (o != ref).nonzero()
        print(f"run {i}: out diffs {len(idx)} comps {idx[:,0].unique().tolist()} samples {idx[:,1].unique().tolist()} nodes {idx[:,3].unique().tolist()} nchan {len(idx[:,2].unique())}")
        ds = (s.view(torch.int32) != sref.view(torch.int32)).nonzero().flatten()
        names = sorted(offs, key=lambda k: offs[k])
        for fi in ds[:: max(1, len(ds) // 12)].tolist():
            m, r = divmod(fi, total)
            nm = [k for k in names if offs[k] <= r][-1]
            print(f"   saved diff comp {m} buffer {nm} rel {r - offs[nm]}")
        print(f"   saved diffs total {len(ds)}")
print(f"{nbad}/120 runs differ")
