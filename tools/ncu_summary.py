#!/usr/bin/env python3
"""Boils an .ncu-rep (ncu --set full) down to the metrics the roofline discussion uses.
usage: ncu_summary.py report.ncu-rep out.json [traffic_out.json]"""
import csv
import json
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "launch__block_size", "launch__grid_size", "launch__registers_per_thread", "launch__cluster_dim_x",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "sm__cycles_active.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__shared_mem_per_block_dynamic"]


def to_bytes(val, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    return float(val) * scale


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    res, traffic = [], []
    for r in rows[2:]:
        d = {"Kernel Name": r[hdr.index("Kernel Name")]}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                d[k] = "%s %s" % (r[i], units[i])
        res.append(d)
        i, j = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        traffic.append(to_bytes(r[i], units[i]) + to_bytes(r[j], units[j]))
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    if len(sys.argv) > 3 and traffic:
        with open(sys.argv[3], "w") as f:
            json.dump({"dram_bytes_per_launch": sum(traffic) / len(traffic), "launches": len(traffic),
                       "kernel": res[0]["Kernel Name"], "source": rep.split("/")[-1]}, f, indent=1)
    print(json.dumps(res[:1], indent=1))


if __name__ == "__main__":
    main()
