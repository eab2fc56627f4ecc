"""Developer helper: tcgen05 / TMEM / TMA / mbarrier opcode counts per kernel of the built library (cuobjdump -sass).

    python tools/sass_histogram.py > profiles/r02_sass_histogram.txt
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "msgat_b200", "csrc", "libmsgat_sm100.so")
OPS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "MUFU.EX2", "FFMA2", "LDS", "STS", "LD.E", "ST.E", "LDG", "STG", "DADD", "DFMA", "REDUX"]
sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
fn, counts, size = None, collections.OrderedDict(), collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[fn] = collections.Counter()
        continue
    if fn is None or "/*" not in line:
        continue
    m = re.search(r"\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", line)
    if not m:
        continue
    op = m.group(1)
    size[fn] += 1
    for o in OPS:
        if op == o or op.startswith(o + "."):
            counts[fn][o] += 1
print(f"{'kernel':44s} {'instr':>6s} " + " ".join(f"{o:>8s}" for o in OPS))
tot = collections.Counter()
for fn, c in counts.items():
    if not any(c[o] for o in ("UTCHMMA", "LDTM", "UBLKCP", "SYNCS")) and size[fn] < 1500:
        continue
    print(f"{fn[:44]:44s} {size[fn]:6d} " + " ".join(f"{c[o]:8d}" for o in OPS))
    tot.update(c)
print(f"{'total (listed kernels)':44s} {'':6s} " + " ".join(f"{tot[o]:8d}" for o in OPS))
