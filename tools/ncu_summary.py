#!/usr/bin/env python
"""Print the key metrics of every kernel in an .ncu-rep (run here, no GPU needed): python tools/ncu_summary.py file.ncu-rep"""
import csv
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("smsp__inst_executed.sum", "winst"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conf"),
]
STALLS = ["long_scoreboard", "barrier", "short_scoreboard", "wait", "lg_throttle", "mio_throttle", "math_pipe_throttle",
          "branch_resolving", "no_instruction", "membar", "not_selected", "dispatch_stall", "drain", "imc_miss", "sleeping"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    for r in data:
        print("==", r[col["Kernel Name"]][:90])
        line = []
        for m, short in WANT:
            if m in col:
                line.append("%s=%s%s" % (short, r[col[m]], units[col[m]] if units[col[m]] not in ("", "%") else ""))
        print("  ", "  ".join(line))
        st = []
        for s in STALLS:
            m = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % s
            if m in col:
                try:
                    st.append((float(r[col[m]]), s))
                except ValueError:
                    pass
        st.sort(reverse=True)
        print("   stalls(warps/issue):", "  ".join("%s=%.2f" % (s, v) for v, s in st[:6]))


if __name__ == "__main__":
    main(sys.argv[1])
