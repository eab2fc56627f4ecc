"""Turn `ncu --set full` reports into the markdown summary kept under profiles/ (developer helper).

    python tools/ncu_summary.py out.md rep1.ncu-rep [rep2.ncu-rep ...]
"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("time", "gpu__time_duration.sum"),
    ("dram read", "dram__bytes_read.sum"),
    ("dram write", "dram__bytes_write.sum"),
    ("regs/thread", "launch__registers_per_thread"),
    ("dyn smem/block", "launch__shared_mem_per_block_dynamic"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
    ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("issue active %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("FMA pipe active %", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("SM throughput %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("DRAM throughput %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor pipe active % (sm__pipe_tensor_cycles_active)", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("tensor pipe active %, elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
    ("tcgen05 instructions % (sm__inst_executed_pipe_tc)", "sm__inst_executed_pipe_tc.sum.pct_of_peak_sustained_active"),
    ("TMEM instructions % (sm__inst_executed_pipe_tmem)", "sm__inst_executed_pipe_tmem.sum.pct_of_peak_sustained_active"),
    ("L2 throughput %", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L2 hit %", "lts__t_sector_hit_rate.pct"),
    ("warp instructions", "smsp__inst_executed.sum"),
    ("stall long scoreboard (samples)", "smsp__pcsamp_warps_issue_stalled_long_scoreboard"),
    ("stall barrier (samples)", "smsp__pcsamp_warps_issue_stalled_barrier"),
    ("stall short scoreboard (samples)", "smsp__pcsamp_warps_issue_stalled_short_scoreboard"),
    ("stall wait (samples)", "smsp__pcsamp_warps_issue_stalled_wait"),
    ("stall not selected (samples)", "smsp__pcsamp_warps_issue_stalled_not_selected"),
    ("stall no instructions (samples)", "smsp__pcsamp_warps_issue_stalled_no_instructions"),
    ("stall math pipe throttle (samples)", "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle"),
    ("stall mio throttle (samples)", "smsp__pcsamp_warps_issue_stalled_mio_throttle"),
    ("smem bank conflicts", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
]


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    dst, reps = sys.argv[1], sys.argv[2:]
    lines = []
    for rep in reps:
        hdr, units, rows = rows_of(rep)
        for r in rows:
            name = r[hdr.index("Kernel Name")]
            lines.append(f"## {name}  ({rep.split('/')[-1]})\n")
            lines.append("| metric | value |\n|---|---|")
            for label, key in METRICS:
                if key in hdr:
                    i = hdr.index(key)
                    lines.append(f"| {label} | {r[i]} {units[i]} |")
            lines.append("")
    with open(dst, "a") as f:
        f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
