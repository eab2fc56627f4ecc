"""Developer helper: which NON-library kernels (torch ops around the hot path) run in a training step."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
sys.argv = [sys.argv[0], "--graph", "off", "--no-cpu-baseline"] + sys.argv[1:]
args = bench.parse()
from msgat_b200 import synthetic
shape = dict(synthetic.PEMS_SHAPES[args.config])
run = bench.Run(args, shape, args.batch or shape["B"], 0, 1, 0)
for _ in range(3):
    run.step(*run.devb)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        run.step(*run.devb)
    torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0 and e.device_type == torch.autograd.DeviceType.CUDA]
rows.sort(key=lambda r: -r[2])
tot = sum(r[2] for r in rows)
print(f"total device time per step: {tot / 3 / 1e3:.3f} ms")
for k, c, t in rows[:45]:
    print(f"{t / 3:9.1f} us/step  x{c / 3:5.1f}  {k[:110]}")
