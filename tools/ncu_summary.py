#!/usr/bin/env python
"""Summarise `ncu --page raw --csv` exports: one block per captured launch with the metrics the roofline uses."""
import csv
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "launch__grid_size", "launch__block_size",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    for row in rows[2:]:
        d, u = dict(zip(hdr, row)), dict(zip(hdr, units))
        print("kernel:", d["Kernel Name"][:110])
        for k in KEYS:
            if k in d:
                print(f"  {k:62s} {d[k]} {u[k]}")
        try:
            rd, wr = float(d["dram__bytes_read.sum"]), float(d["dram__bytes_write.sum"])
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
            tb = rd * scale[u["dram__bytes_read.sum"]] + wr * scale[u["dram__bytes_write.sum"]]
            t = float(d["gpu__time_duration.sum"]) * {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1}[u["gpu__time_duration.sum"]]
            print(f"  {'dram traffic (read+write)':62s} {tb / 1e9:.4f} GB  -> {tb / t / 1e9:.0f} GB/s")
        except Exception:
            pass
