#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed) into a small text table for profiles/.

  python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_<what>.txt
"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__maximum_warps_per_active_cycle_pct",
    "launch__registers_per_thread",
    "launch__grid_size",
    "launch__block_size",
    "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors.sum",
    "lts__t_sector_hit_rate.pct",
    "sm__cycles_elapsed.max",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE,
                         stderr=subprocess.DEVNULL).stdout.decode(errors="replace")
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    print("# %s (ncu --set full --clock-control none; per launch)" % rep)
    for r in rows[2:]:
        print("\n== %s  grid %s block %s" % (r[col["Kernel Name"]][:110], r[col.get("Grid Size", 0)],
                                             r[col.get("Block Size", 0)]))
        for w in WANT:
            if w in col:
                print("  %-66s %16s %s" % (w, r[col[w]], units[col[w]]))
        if "dram__bytes_read.sum" in col:
            try:
                tot = float(r[col["dram__bytes_read.sum"]]) + float(r[col["dram__bytes_write.sum"]])
                print("  %-66s %16.3f %s" % ("traffic = dram read + write", tot, units[col["dram__bytes_read.sum"]]))
            except ValueError:
                pass


if __name__ == "__main__":
    main()
