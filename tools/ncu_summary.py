#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i ... --page raw --csv`) into one line per launch."""
import csv, subprocess, sys
WANT = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"),
        ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("smsp__inst_executed.sum", "inst"),
        ("launch__grid_size", "grid"), ("launch__block_size", "blk")]
def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    cols = [(h.index(k), n) for k, n in WANT if k in h]
    print(" | ".join(n + ("[" + units[i] + "]" if units[i] else "") for i, n in cols))
    for r in rows[2:]:
        print(" | ".join(r[i].split("(")[0][:28] for i, n in cols))
if __name__ == "__main__":
    main(sys.argv[1])
