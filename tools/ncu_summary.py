#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the handful of numbers DESIGN.md and
profiles/ quote.  Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__inst_executed.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct",
        "l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_ld.ratio", "l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_st.ratio"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        print("==", r[ki][:70])
        for k in KEYS:
            if k in hdr:
                print(f"   {k:85s} {r[hdr.index(k)]:>16s} {units[hdr.index(k)]}")
        st = [(float(r[i]), h) for i, h in enumerate(hdr) if "average_warps_issue_stalled" in h and "per_issue_active" in h and r[i] not in ("", "n/a")]
        for v, h in sorted(st, reverse=True)[:6]:
            print(f"   stall {h.split('issue_stalled_')[1].split('_per_issue')[0]:30s} {v:6.2f}")


if __name__ == "__main__":
    main(sys.argv[1])
