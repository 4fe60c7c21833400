#!/usr/bin/env python
"""Compact per-kernel table from an .ncu-rep (ncu -i ... --page raw --csv).  Usage: ncu_summary.py file.ncu-rep"""
import csv
import io
import subprocess
import sys

COLS = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"),
        ("dram__bytes_write.sum", "wrMB"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "fmaheavy%"),
        ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed", "alu%"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "fp64%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed", "lsu%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "winst"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bankconf"),
        ("lts__t_bytes.sum", "L2MB")]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print("| " + " | ".join(c[1] for c in COLS) + " |")
    print("|" + "---|" * len(COLS))
    for r in rows[2:]:
        cells = []
        for name, short in COLS:
            i = idx.get(name)
            v = r[i] if i is not None else ""
            if short == "kernel":
                v = v.split("(")[0].replace("shelfi::", "")
            else:
                try:
                    f = float(v.replace(",", ""))
                    u = units[i]
                    if short in ("rdMB", "wrMB", "L2MB"):
                        f = f * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1)
                    if short == "us":
                        f = f * {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "msecond": 1e3, "nsecond": 1e-3}.get(u, 1)
                    v = f"{f:.1f}" if abs(f) < 1e6 else f"{f:.3g}"
                except ValueError:
                    pass
            cells.append(v)
        print("| " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
