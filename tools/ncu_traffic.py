#!/usr/bin/env python
"""profiles/traffic*.json from the raw-page CSV of an `ncu --set full` capture of one wave (tools/ncu_capture.sh): per kernel
class, the average DRAM bytes (read + write) per launch and the duration-weighted issue / lane / occupancy / cache figures.
Usage: python tools/ncu_traffic.py out.json raw.csv [label]"""
import csv
import json
import re
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "sector": 32.0}
TIME = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3, "second": 1e6}


def cls_of(name):
    for c in ("k_extend_pt", "k_extend", "k_shade", "k_connect"):
        if re.search(r"\b" + c + r"\b", name):
            return c
    return None


def main(out, path, label=""):
    rows = list(csv.reader(open(path)))
    while rows and (not rows[0] or rows[0][0] != "ID"):
        rows.pop(0)
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    kn = col["Kernel Name"]

    def num(r, m):
        i = col[m]
        v = float(r[i].replace(",", "")) if r[i] not in ("", "n/a") else 0.0
        return v * UNIT.get(units[i], 1.0)

    def us(r):
        i = col["gpu__time_duration.sum"]
        return float(r[i].replace(",", "")) * TIME.get(units[i], 1.0)
    res = {}
    for c in ("k_extend_pt", "k_extend", "k_shade", "k_connect"):
        rs = [r for r in body if cls_of(r[kn]) == c]
        if not rs:
            continue
        tt = sum(us(r) for r in rs)

        def wmean(m):
            return sum(num(r, m) * us(r) for r in rs) / tt if m in col and tt > 0 else None
        rd, wr = sum(num(r, "dram__bytes_read.sum") for r in rs), sum(num(r, "dram__bytes_write.sum") for r in rs)
        l2rd = sum(num(r, "lts__t_sectors_op_read.sum") for r in rs) if "lts__t_sectors_op_read.sum" in col else None
        e = {"launches": len(rs), "gpu_time_us_per_launch": tt / len(rs), "gpu_time_us_total": tt,
             "dram_bytes_per_launch": (rd + wr) / len(rs), "dram_read_per_launch": rd / len(rs), "dram_write_per_launch": wr / len(rs),
             "dram_read_gbs": rd / tt / 1e3, "dram_gbs": (rd + wr) / tt / 1e3,
             "issue_active_pct": wmean("smsp__issue_active.avg.pct_of_peak_sustained_active"),
             "threads_per_instruction": wmean("smsp__thread_inst_executed_per_inst_executed.ratio"),
             "warps_active_pct": wmean("sm__warps_active.avg.pct_of_peak_sustained_active"),
             "dram_throughput_pct": wmean("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
             "l1_hit_pct": wmean("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": wmean("lts__t_sector_hit_rate.pct"),
             "registers": wmean("launch__registers_per_thread"),
             "source": f"ncu --set full --clock-control none, one wave (2^22 paths){' of ' + label if label else ''}: {len(rs)} launches, raw page {path.split('/')[-1]}"}
        if l2rd is not None:
            e["l2_read_gbs"] = l2rd / tt / 1e3
        res[c] = {k: v for k, v in e.items() if v is not None}
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    for c, e in res.items():
        print(c, {k: (round(v, 2) if isinstance(v, float) else v) for k, v in e.items() if k != "source"})


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
