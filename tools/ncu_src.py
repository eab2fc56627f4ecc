"""Developer helper: top stall lines of an ncu report (source page) + a few raw metrics."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tc.sum", "sm__inst_executed_pipe_tmem.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
for r in rows[2:]:
    print("kernel:", r[hdr.index("Kernel Name")][:60])
    for w in want:
        for i, h in enumerate(hdr):
            if h.startswith(w):
                print(f"  {h[:70]:70s} {r[i]} {rows[1][i]}")
                break
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"] if False else ["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
# find header
hi = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
hdr = rows[hi]; data = rows[hi + 1:]
si = hdr.index("# Samples"); so = hdr.index("Source"); ie = hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[si] or 0) for r in data if len(r) > si)
print("total samples", tot)
agg = collections.Counter()
for r in data:
    if len(r) <= si: continue
    for i in stall_cols:
        agg[hdr[i]] += int(r[i] or 0)
print("stall totals:", ", ".join(f"{k}={v}" for k, v in agg.most_common(8)))
top = sorted(range(len(data)), key=lambda i: -int(data[i][si] or 0))[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]
for i in sorted(top):
    r = data[i]
    st = sorted(((int(r[c] or 0), hdr[c]) for c in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {r[si]:>6s} {r[ie]:>9s}  {r[so].strip()[:70]:70s} {st}")
