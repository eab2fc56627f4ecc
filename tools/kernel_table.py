"""Developer helper: the per-kernel markdown table of DESIGN.md section 4 from a bench.py JSON line.

    python tools/kernel_table.py profiles/r02_bench_final.json [peak_GBps]
"""
import json
import sys

d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
peak = float(sys.argv[2]) if len(sys.argv) > 2 else d["roofline"]["peak"]
print(f"{d['value']:.0f} samples/s, {d['ms_per_step']:.2f} ms/step, e2e {d['e2e']['value']:.0f} samples/s, "
      f"{d['gpu_launches']} launches/step, sum of kernel times {d['roofline']['my_kernels_ms_per_step']:.2f} ms\n")
print("| Kernel (PEMSD7, B=32) | launches/step | us/launch | algorithmic GB/s | fraction of HBM peak | share of device time |")
print("|---|---|---|---|---|---|")
for k, v in sorted(d["kernels"].items(), key=lambda kv: -kv[1]["share"]):
    g = v.get("GBps")
    print(f"| `{k}` | {v['launches_per_step']:.0f} | {v['ms_per_launch'] * 1e3:.0f} | {'' if g is None else f'{g:.0f}'} | "
          f"{'' if g is None else f'{g / peak:.2f}'} | {v['share'] * 100:.1f} % |")
