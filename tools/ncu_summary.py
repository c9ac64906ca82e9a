#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): one record per captured launch with the metrics the
roofline discussion uses.   python tools/ncu_summary.py <report.ncu-rep> [labels,comma,separated] > summary.json"""
import csv
import io
import json
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
    "sm__cycles_elapsed.avg.per_second", "dram__cycles_elapsed.avg.per_second",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "lts__t_sectors_srcunit_tex_aperture_peer_op_read.sum", "lts__t_sectors_srcunit_tex_aperture_sysmem_op_write.sum",
]


def main():
    rep = sys.argv[1]
    labels = sys.argv[2].split(",") if len(sys.argv) > 2 else []
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_col = hdr.index("Kernel Name")
    launches = []
    for i, r in enumerate(data):
        rec = {"label": labels[i] if i < len(labels) else None, "Kernel Name": r[name_col]}
        for k in KEEP:
            if k in hdr:
                j = hdr.index(k)
                rec[k] = f"{r[j]} {units[j]}".strip()
        try:
            rd = float(r[hdr.index("dram__bytes_read.sum")]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[hdr.index("dram__bytes_read.sum")]]
            wr = float(r[hdr.index("dram__bytes_write.sum")]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[hdr.index("dram__bytes_write.sum")]]
            rec["dram_traffic_bytes"] = round(rd + wr)
        except Exception:
            pass
        launches.append(rec)
    print(json.dumps({"report": rep, "launches": launches}, indent=1))


if __name__ == "__main__":
    main()
