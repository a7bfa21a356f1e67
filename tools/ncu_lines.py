#!/usr/bin/env python
"""Per-source-line cost of one kernel: joins the SASS page of an .ncu-rep (executed warp instructions and
stall samples per instruction) with `nvdisasm -g` line info of the SAME binary.
Usage: python tools/ncu_lines.py file.ncu-rep libplugpu.so <mangled-kernel-substring> [block-index] [top-n]"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile


def sass_rows(rep, which):
    """rep: an .ncu-rep, or the gzipped `--page source --csv --print-source sass` export tools/ncu_capture.sh leaves; which: block
    index, or a kernel-name substring (then the launch of that kernel with the most samples)"""
    if rep.endswith(".gz"):
        import gzip
        out = gzip.open(rep, "rt").read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    blocks, cur = [], None
    for row in csv.reader(out.splitlines()):
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "rows": [], "hdr": None}; blocks.append(cur); continue
        if cur is None or not row:
            continue
        if cur["hdr"] is None:
            cur["hdr"] = row; continue
        cur["rows"].append(row)
    if isinstance(which, str):
        cand = [b for b in blocks if which in b["name"]]
        return max(cand, key=lambda b: sum(int(r[b["hdr"].index("# Samples")] or 0) for r in b["rows"]))
    return blocks[which]


def line_table(so, kernel, outer=False):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    lines = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        dis = subprocess.run(["nvdisasm", "-gi" if outer else "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        inside, cur, fresh = False, ("?", 0), True
        for l in dis.splitlines():
            if l.startswith("//---") and ".text." in l:
                inside = kernel in l
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
            if m:
                if outer and m.group(3):   # the chain ends at the outermost call site
                    cur = (os.path.basename(m.group(3)), int(m.group(4)))
                elif not outer or fresh:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                fresh = False
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
                lines.append(cur); fresh = True
    return lines


def main(rep, so, kernel, which=0, top=40, outer=False):
    b = sass_rows(rep, which)
    lt = line_table(so, kernel, outer)
    h = b["hdr"]
    iN, iE = h.index("# Samples"), h.index("Instructions Executed")
    print(b["name"][:90], "| sass rows", len(b["rows"]), "| disasm rows", len(lt))
    n = min(len(lt), len(b["rows"]))
    samp, exe = collections.Counter(), collections.Counter()
    for i in range(n):
        samp[lt[i]] += int(b["rows"][i][iN] or 0); exe[lt[i]] += int(b["rows"][i][iE] or 0)
    ts, te = sum(samp.values()), sum(exe.values())
    print("by file: ", {f: (round(sum(v for k, v in samp.items() if k[0] == f) / ts, 3), round(sum(v for k, v in exe.items() if k[0] == f) / te, 3)) for f in {k[0] for k in samp}})
    print(f"{'file:line':28s} samples  exec")
    for k, v in samp.most_common(top):
        print(f"{k[0] + ':' + str(k[1]):28s} {v / ts:6.3f} {exe[k] / te:6.3f}")


if __name__ == "__main__":
    a = [x for x in sys.argv if x != "--outer"]   # --outer: attribute inlined code to the outermost call site
    which = (int(a[4]) if a[4].lstrip("-").isdigit() else a[4]) if len(a) > 4 else 0
    main(a[1], a[2], a[3], which, int(a[5]) if len(a) > 5 else 40, "--outer" in sys.argv)
