#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the per-kernel table kept under profiles/."""
import csv
import sys

WANT = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "ms"),
        ("dram__bytes_read.sum", "dram_read_GB"), ("dram__bytes_write.sum", "dram_write_GB"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64_pipe_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
        ("launch__registers_per_thread", "regs"), ("launch__occupancy_limit_shared_mem", "cta_limit_smem"),
        ("launch__occupancy_limit_registers", "cta_limit_regs"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall_barrier"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall_long_sb"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall_short_sb"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall_math_throttle"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall_wait"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts")]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    cols = [(hdr.index(k), name) for k, name in WANT if k in hdr]
    out = csv.writer(sys.stdout)
    out.writerow([name for _, name in cols])
    for r in rows[2:]:
        vals = []
        for i, name in cols:
            v = r[i]
            if name == "kernel":
                v = v.split("(")[0].replace("void ", "")
            elif name == "ms" and units[i] in ("us", "usecond"):
                v = "%.4f" % (float(v) / 1e3)
            elif name == "ms" and units[i] in ("ns", "nsecond"):
                v = "%.4f" % (float(v) / 1e6)
            elif name.endswith("_GB"):
                scale = {"Gbyte": 1.0, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}.get(units[i], 1.0)
                v = "%.4f" % (float(v) * scale)
            else:
                try:
                    v = "%.3f" % float(v) if "." in v else v
                except ValueError:
                    pass
            vals.append(v)
        out.writerow(vals)


if __name__ == "__main__":
    main(sys.argv[1])
