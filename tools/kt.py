"""Print the per-kernel table of a bench.py JSON line (developer helper): python tools/kt.py gpurun_out/x.json [filter]"""
import json
import signal
import sys

signal.signal(signal.SIGPIPE, signal.SIG_DFL)   # `| head` closes the pipe early

d = json.load(open(sys.argv[1]))
flt = sys.argv[2] if len(sys.argv) > 2 else ""
print(f"{d['ms_per_step']:.3f} ms/step  {d['value']:.0f} {d['unit']}  e2e {d['e2e']['ms_per_step']:.3f} ms")
for k, v in sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_launch"] * kv[1]["launches_per_step"]):
    if flt in k:
        print(f"{k:28s} {v['ms_per_launch'] * 1000:8.1f} us {v['share'] * 100:5.1f}%")
