#!/usr/bin/env python
"""Hot spots of one kernel from an .ncu-rep source page: stall samples by SASS instruction and opcode.
Usage: python tools/ncu_hot.py file.ncu-rep [kernel-index] [top-n]"""
import collections
import csv
import subprocess
import sys


def main(path, which=0, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    blocks, cur = [], None
    for row in csv.reader(out.splitlines()):
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "rows": [], "hdr": None}; blocks.append(cur); continue
        if cur is None or not row:
            continue
        if cur["hdr"] is None:
            cur["hdr"] = row; continue
        cur["rows"].append(row)
    b = blocks[which]
    h = b["hdr"]
    iS, iN, iE, iT = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed"), h.index("Avg. Threads Executed")
    stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[iN] or 0) for r in b["rows"])
    tot_inst = sum(int(r[iE] or 0) for r in b["rows"])
    print(b["name"][:100], "samples", tot, "warp-instr", tot_inst, "sass lines", len(b["rows"]))
    by_op, inst_op = collections.Counter(), collections.Counter()
    reasons = collections.Counter()
    for r in b["rows"]:
        op = r[iS].split()[0] if not r[iS].strip().startswith("@") else r[iS].split()[1]
        op = op.split(".")[0]
        by_op[op] += int(r[iN] or 0); inst_op[op] += int(r[iE] or 0)
        for i in stall_cols:
            reasons[h[i]] += int(r[i] or 0)
    print("stall reasons:", [(k, round(v / max(tot, 1), 3)) for k, v in reasons.most_common(8)])
    print("by opcode (share of samples | share of executed warp-instr):")
    for op, n in by_op.most_common(18):
        print(f"   {op:10s} {n / max(tot, 1):6.3f} {inst_op[op] / max(tot_inst, 1):6.3f}")
    print("hottest instructions:")
    rows = sorted(enumerate(b["rows"]), key=lambda x: -int(x[1][iN] or 0))[:top]
    for idx, r in sorted(rows):
        why = max(stall_cols, key=lambda i: int(r[i] or 0))
        print(f"   #{idx:5d} {int(r[iN] or 0) / max(tot, 1):6.3f} exec {r[iE]:>9s} thr {r[iT]:>5s} {h[why]:18s} {r[iS].strip()[:90]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else 40)
