#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python tools/launch_summary.py launches.csv > summary.txt"""
import collections
import csv
import re
import sys


def main(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv, im = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows[1:]:
        if r[im] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("plud::", "").replace("(bool)", "").replace("(int)", "")
        tot[name] += float(r[iv].replace(",", "")) / 1e3; cnt[name] += 1
    all_us = sum(tot.values())
    print(f"# {path}: {sum(cnt.values())} launches, {all_us:.1f} us; per-launch times are cold-cache and serialised: compare SHARES")
    print(f"{'kernel':60s} {'launches':>8s} {'total_us':>12s} {'share':>7s}")
    for k, v in tot.most_common():
        print(f"{k[:60]:60s} {cnt[k]:8d} {v:12.1f} {v / all_us:7.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
