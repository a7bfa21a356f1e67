#!/usr/bin/env python
"""profiles/traffic.json from `ncu --set full` captures: average DRAM bytes (read + write) per launch of each
kernel class over the captured launches (capture whole waves: 8 k_extend, 16 k_shade, 8 k_connect each).
Usage: python tools/ncu_traffic.py out.json class=file.ncu-rep [class=file.ncu-rep ...]"""
import csv
import json
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(out, pairs):
    res = {}
    for pair in pairs:
        cls, path = pair.split("=", 1)
        txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(txt.splitlines()))
        hdr, units = rows[0], rows[1]
        ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
        rd = sum(float(r[ir]) * UNIT[units[ir]] for r in rows[2:])
        wr = sum(float(r[iw]) * UNIT[units[iw]] for r in rows[2:])
        n = len(rows) - 2

        def mean(metric):   # launch-duration-weighted mean of a per-launch percentage / ratio
            i = hdr.index(metric)
            tt = sum(float(r[it]) for r in rows[2:])
            return sum(float(r[i]) * float(r[it]) for r in rows[2:]) / tt
        res[cls] = {}
        res[cls].update({"issue_active_pct": mean("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "threads_per_instruction": mean("smsp__thread_inst_executed_per_inst_executed.ratio"),
                    "warps_active_pct": mean("sm__warps_active.avg.pct_of_peak_sustained_active"),
                    "dram_throughput_pct": mean("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")})
        res[cls].update({"dram_bytes_per_launch": (rd + wr) / n, "dram_read_per_launch": rd / n, "dram_write_per_launch": wr / n,
                    "launches": n, "gpu_time_us_per_launch": sum(float(r[it]) for r in rows[2:]) / n,
                    "source": f"ncu --set full --clock-control none, {n} consecutive launches (whole waves), capture {path.split('/')[-1]} (headline metrics in profiles/r01_final_ncu_summary.txt)"})
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
