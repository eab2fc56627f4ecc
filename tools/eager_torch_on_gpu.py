"""Informational "before" number (SURVEY.md 8c variants iii / iv): the reference's algorithm as stock
PyTorch eager ops on the B200 -- the oracle port (oracle/msgat_oracle.py) with its tensors on cuda:0, fp32
(allow_tf32 off) and under autocast as src/engine.py:54 runs it.  Not a bench arm (the reference arm of
bench.py is the CPU path); prints one JSON line per precision.  Developer helper, run through gpurun.

    python tools/eager_torch_on_gpu.py [--config pemsd7] [--steps 10]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="pemsd7")
    ap.add_argument("--model", default="msgat72")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    from msgat_b200 import synthetic
    from oracle import msgat_oracle as O
    shape = synthetic.PEMS_SHAPES[args.config]
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    adj = synthetic.make_adjacency(args.config)
    p = {k: v.to(dev) for k, v in O.init_params(args.model, shape["M"], shape["C"], adj).items()}
    train = [v.requires_grad_(True) for k, v in p.items() if k != "adj"]
    opt = torch.optim.Adam(train, lr=1e-3, weight_decay=5e-4)
    batch = synthetic.make_batch(shape["B"], shape["M"], shape["C"], shape["N"], seed=1234)
    X, H, D, Y = (t.to(dev) for t in (batch.X, batch.H, batch.D, batch.Y))
    for mode in ("fp32", "autocast"):
        def step():
            with torch.autocast("cuda", enabled=(mode == "autocast")):
                loss = O.train_step_loss(p, args.model, X, H, D, Y, 50.0)
            opt.zero_grad()
            loss.backward()
            opt.step()
        for _ in range(args.warmup):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        print(json.dumps({"what": "oracle port as stock PyTorch eager ops on cuda:0 (not a bench arm)", "precision": mode,
                          "config": args.config, "batch": shape["B"], "ms_per_step": ms,
                          "samples_per_s": shape["B"] / ms * 1e3,
                          "peak_mem_GB": torch.cuda.max_memory_allocated() / 2 ** 30}))


if __name__ == "__main__":
    main()
