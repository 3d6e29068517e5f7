#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed) into the few counters the roofline uses.
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_xxx.txt"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum.per_second",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.avg", "sm__cycles_active.avg", "smsp__cycles_active.avg",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.avg.per_second"]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    names = [i for i, h in enumerate(hdr) if h == "Kernel Name"]
    print("# ncu --set full --clock-control none summary of %s" % rep)
    for r in rows[2:]:
        print("kernel: %s" % (r[names[0]] if names else "?"))
        for w in WANT:
            for i, h in enumerate(hdr):
                if h == w:
                    print("  %-70s %s %s" % (h, r[i], units[i]))
        rd = [float(r[i]) for i, h in enumerate(hdr) if h == "dram__bytes_read.sum"]
        un = [units[i] for i, h in enumerate(hdr) if h == "dram__bytes_read.sum"]
        if rd:
            print("  (traffic = dram read + write, units as printed: %s %s)" % (rd[0], un[0]))


if __name__ == "__main__":
    main()
