"""Developer helper: warp-state samples of an ncu report aggregated per CUDA source line.
usage: ncu_lines.py <report.ncu-rep> <object.o> <kernel-mangled-substring> [top]
The SASS rows of `ncu --page source` are matched, in order, with the instructions of `nvdisasm -g` of the same
object (built with -lineinfo), whose `//## File ..., line N` markers give the line of each instruction (the innermost
inlined location)."""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, obj, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
outer = os.environ.get("OUTER") == "1"   # attribute to the outermost call site (the kernel's own line) instead of the inlined helper
dis = subprocess.run(["nvdisasm", "-gi" if outer else "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
lines, cur, infn, stack = [], None, False, None
for l in dis.splitlines():
    if l.startswith("\t.section\t.text."):
        infn = kname in l
        continue
    if not infn:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))   # with -gi the last marker of a group is the outermost frame
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l):
        lines.append(cur)
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
his = [i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r]
which = int(os.environ.get("KERNEL_INDEX", "0"))   # which launch of a multi-launch report
hi = his[which]
end = his[which + 1] if which + 1 < len(his) else len(rows)
hdr, data = rows[hi], [r for r in rows[hi + 1:end] if len(r) > 3 and r[0].startswith("0x")]
si = hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
print(f"sass rows {len(data)}, disassembled instructions {len(lines)}")
agg, why = collections.Counter(), collections.defaultdict(collections.Counter)
for i, r in enumerate(data):
    key = lines[i] if i < len(lines) else ("?", 0)
    n = int(r[si] or 0)
    agg[key] += n
    for c in stall:
        why[key][hdr[c]] += int(r[c] or 0)
tot = sum(agg.values())
text = {}
for (f, n), s in agg.most_common(top):
    if f not in text:
        p = os.path.join(os.path.dirname(os.path.abspath(obj)), f)
        text[f] = open(p).read().splitlines() if os.path.exists(p) else []
    t = text[f][n - 1].strip()[:90] if 0 < n <= len(text[f]) else ""
    w = ", ".join(f"{k[6:]}={v}" for k, v in why[(f, n)].most_common(2))
    print(f"{s:7d} {100.0 * s / tot:5.1f}%  {f}:{n:<5d} {t}   [{w}]")
