"""Developer helper: stress the layer-2 grouped forward (k_l2_mix, k_small, score pass, k_l2_out) at the bench shape:
N launches with blocking launches, bit-compares every output with the first one, reports failures (a trap shows as an
exception).  Usage: [env switches] python tools/stress_fwd.py [iterations]"""
import os
import sys
import time
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
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
with torch.no_grad():
    ref = ops.meam_group_op(x, "mb", cfg, sp, [mm.slots() for mm in meams]).clone()
    torch.cuda.synchronize()
    t0, bad = time.time(), 0
    try:
        for i in range(iters):
            out = ops.meam_group_op(x, "mb", cfg, sp, [mm.slots() for mm in meams])
            if i % 16 == 0:
                torch.cuda.synchronize()
                if not torch.equal(out, ref):
                    bad += 1
        torch.cuda.synchronize()
        print(f"ok: {iters} forwards, {bad} mismatching of {(iters + 15) // 16} compared, {time.time() - t0:.1f}s", {k: v for k, v in os.environ.items() if k.startswith("MSGAT")})
    except Exception as e:
        print(f"FAILED after <= {i} forwards: {str(e)[:120]}", {k: v for k, v in os.environ.items() if k.startswith("MSGAT")})
