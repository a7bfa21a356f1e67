#!/usr/bin/env python
"""Stall reasons and opcode hot spots per kernel from a gzipped `ncu --page source --csv --print-source sass` export
(tools/ncu_capture.sh).  The largest launch of each kernel is summarised.
Usage: python tools/ncu_source_summary.py file_source.csv.gz [kernel-substring ...]"""
import collections
import csv
import gzip
import io
import sys


def main(path, want):
    rows = csv.reader(io.TextIOWrapper(gzip.open(path), newline=""))
    blocks, cur = [], None
    for row in rows:
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "rows": [], "hdr": None}; blocks.append(cur); continue
        if cur is None or not row:
            continue
        if cur["hdr"] is None:
            cur["hdr"] = row; continue
        cur["rows"].append(row)
    best = {}
    for b in blocks:
        key = b["name"].split("(plud")[0]
        if want and not any(w in key for w in want):
            continue
        h = b["hdr"]
        iN = h.index("# Samples")
        tot = sum(int(r[iN] or 0) for r in b["rows"])
        if key not in best or tot > best[key][0]:
            best[key] = (tot, b)
    for key, (tot, b) in sorted(best.items()):
        h = b["hdr"]
        iS, iN, iE, iT = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed"), h.index("Avg. Threads Executed")
        stall = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
        ti = sum(int(r[iE] or 0) for r in b["rows"])
        thr = sum(int(r[iE] or 0) * float(r[iT] or 0) for r in b["rows"]) / max(ti, 1)
        reasons, byop, iop = collections.Counter(), collections.Counter(), collections.Counter()
        for r in b["rows"]:
            parts = r[iS].strip().split()
            op = (parts[1] if parts and parts[0].startswith("@") and len(parts) > 1 else (parts[0] if parts else "?")).split(".")[0]
            byop[op] += int(r[iN] or 0); iop[op] += int(r[iE] or 0)
            for i in stall:
                reasons[h[i]] += int(r[i] or 0)
        print(f"== {key.replace('void ', '')}: {tot} samples, {ti} warp instructions, {thr:.1f} threads / instruction")
        print("   stall reasons:", ", ".join(f"{k.replace('stall_', '')} {v / max(tot, 1):.2f}" for k, v in reasons.most_common(7)))
        print("   opcode: share of stall samples / share of executed instructions")
        print("     " + ", ".join(f"{op} {n / max(tot, 1):.2f}/{iop[op] / max(ti, 1):.2f}" for op, n in byop.most_common(10)))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
