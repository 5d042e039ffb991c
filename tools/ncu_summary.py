#!/usr/bin/env python3
"""Turn an .ncu-rep (ncu --set full) into the small per-launch CSV committed under profiles/.

usage: tools/ncu_summary.py gpurun_out/prof_X.ncu-rep profiles/prof_X.csv
"""
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__cycles_elapsed.max",
    "smsp__cycles_active.avg",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True, check=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    head, units = rows[0], rows[1]
    cols = [m for m in METRICS if m in head]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel"] + ["%s [%s]" % (m, units[head.index(m)]) for m in cols])
        for r in rows[2:]:
            name = r[head.index("Kernel Name")]
            name = name.replace("void <unnamed>::", "").replace("<unnamed>::", "")
            w.writerow([r[head.index("ID")], name[:80]] + [r[head.index(m)] for m in cols])
    print("wrote", out, len(rows) - 2, "launches")


if __name__ == "__main__":
    main()
