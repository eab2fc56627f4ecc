"""Developer helper: profiles/ncu_traffic.json (DRAM bytes per launch of each kernel, read by bench.py for roofline.traffic)
from the ncu --set full reports of the bench step.

    python tools/ncu_traffic.py gpurun_out/r02_*.ncu-rep
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
# bench.py's kernel keys (name@C..O..) of the PEMSD7 workload for the captured launches
KEYS = {"k_tc_dgrad": "k_tc_dgrad@C72O72", "k_tc_wgrad": "k_tc_wgrad@C72O72", "k_l2_out": "k_l2_out<24>@C72O72",
        "k_l2_mix": "k_l2_mix<24>@C72O72", "k_l2_gz": "k_l2_gz@C72O72", "k_head_bwd12": "k_head_bwd12@C72O12",
        "k_head_fwd12": "k_head_fwd12@C72O12", "k_bwd_fix": "k_bwd_fix<16>@C72O72", "k_l1_bwd": "k_l1_bwd<1>@C1O72",
        "k_l1_fwd": "k_l1_fwd<1>@C1O72", "k_score_tc32<0>": "k_score_tc32<0>@C72O72", "k_score_tc32<1>": "k_score_tc32<1>@C72O72",
        "k_score_tc<0>": "k_score_tc<0>@C72O72", "k_score_tc<1>": "k_score_tc<1>@C72O72", "k_sstats": "k_sstats@C72O72",
        "k_scol": "k_scol@C72O72"}


def main():
    out = {"_source": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full --clock-control none, PEMSD7 shape B=32 "
                      "(round 2; summaries in profiles/r02_ncu_top_kernels.md)", "pemsd7:fp32": {}, "pemsd7:tf32": {}}
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units = rows[0], rows[1]
        for r in rows[2:]:
            name = r[hdr.index("Kernel Name")]
            tot = 0.0
            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                i = hdr.index(m)
                tot += float(r[i]) * UNIT[units[i]]
            m = re.match(r"(?:void )?(k_[a-z0-9_]+(?:<[0-9]>)?)", name)
            base = m.group(1) if m else name
            key = KEYS.get(base) or KEYS.get(re.sub(r"<.*", "", base))
            if key is None:
                continue
            tf32_only = base.startswith("k_score_tc<")
            if not tf32_only:
                out["pemsd7:fp32"][key] = int(tot)
            if not base.startswith(("k_score_tc32", "k_sstats", "k_scol")):
                out["pemsd7:tf32"][key] = int(tot)
    with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
