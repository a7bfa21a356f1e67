#!/usr/bin/env python
"""profiles/r02_sass_summary.txt: per kernel of libplugpu.so (sm_100a) — registers, stack, spills (cuobjdump -res-usage) and the
counts of the SASS mnemonics that characterise the path (cuobjdump -sass):
  FMNMX3 / FMNMX.NAN  three-input and NaN-propagating min/max of the slab tests (isect.cuh: slabs_fast)
  FADD2 / FMUL2       packed fp32x2 arithmetic of the fast-BVH node test (two children per issue slot)
  LDGSTS              cp.async gathers of the staged loops (wavefront.cuh)
  MUFU / FCHK         reciprocal / rsqrt seeds and range checks of the IEEE divisions and square roots (exact flags)
  STL / LDL           local-memory traffic (traversal stack levels beyond the shared-memory ones, spills)
  CALL                slow paths of div / sqrt
No UTMALDG / UTCMMA / tcgen05 are expected: the path is not a dense contraction.
Usage: python tools/sass_summary.py [lib.so] > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WATCH = ["FMNMX3", "FMNMX.NAN", "FADD2", "FMUL2", "FFMA", "LDGSTS", "MUFU", "FCHK", "CALL", "STL", "LDL", "LDG", "STG", "ATOM", "RED", "BAR", "UTMALDG", "UTCMMA"]


def demangle(names):
    out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main(lib):
    res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1); continue
        if cur and "REG:" in line:
            usage[cur] = dict(re.findall(r"(\w+):(\d+)", line)); cur = None
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    counts, total, arch = collections.defaultdict(collections.Counter), collections.Counter(), set()
    fn = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = m.group(1); continue
        m = re.match(r"\s*arch = (\S+)", line)
        if m:
            arch.add(m.group(1)); continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", line)
        if m and fn:
            op = m.group(1)
            total[fn] += 1
            base = op.split(".")[0]
            counts[fn][base] += 1
            if op.startswith("FMNMX") and ".NAN" in op:
                counts[fn]["FMNMX.NAN"] += 1
    names = demangle(sorted(total))
    print(f"# {os.path.relpath(lib, ROOT)}  arch: {', '.join(sorted(arch))}   (tools/sass_summary.py)")
    print("# kernel | regs | stack | SASS instr | " + " | ".join(WATCH))
    for fn in sorted(total, key=lambda f: names[f]):
        if not re.search(r"plud::|plubuild::", names[fn]):
            continue
        u = usage.get(fn, {})
        short = re.sub(r"\(.*", "", names[fn]).replace("void ", "")
        print(f"{short} | {u.get('REG', '?')} | {u.get('STACK', '?')} | {total[fn]} | " + " | ".join(str(counts[fn].get(w, 0)) for w in WATCH))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "plutracer_b200", "_lib", "libplugpu.so"))
