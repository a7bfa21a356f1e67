#!/usr/bin/env python
"""tests/golden/lattice.pluscene: a scene whose box centres TIE along every axis, loaded and built by the
UNMODIFIED reference (oracle/_ref/plu_ref dump-scene), so that the leaf ranks in the file are what libstdc++'s
unstable std::sort left in the reference's live bvh_tree (bvh_tree.cpp:23-27) — the case in which a restatement
of the build must reproduce introsort's order of equal elements, not just "a" sorted order.

    python tests/golden/make_lattice_golden.py        (build container only: needs /root/reference)

12^3 unit boxes on the integer lattice (144 equal centres per coordinate value and axis) followed by 60 copies of
one box (identical centres on all axes), a point light and an area light."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "plu_ref")


def main():
    urn = "/tmp/plu_lattice.urn"
    with open(urn, "w") as f:
        f.write("[\n\tresolution: [64 64]\n\tantialiasing-samples: 2\n\tcamera: [ position: [0 0 -40] target: [0 0 0] up: [0 1 0] ]\n\tmaterials: [\n")
        for k, c in enumerate(("0.8 0.3 0.3", "0.3 0.8 0.3", "0.3 0.3 0.8", "0.7 0.7 0.7")):
            f.write(f"\t\tm{k}: [ diffuse [{c}] ]\n")
        f.write("\t]\n\tobjects: [\n")
        i = 0
        for x in range(12):
            for y in range(12):
                for z in range(12):
                    f.write(f"\t\tbox [{2 * x - 11} {2 * y - 11} {2 * z - 11}] [0.5 0.5 0.5] 'm{i % 4}\n")
                    i += 1
        for k in range(60):
            f.write(f"\t\tbox [0 14 0] [0.75 0.25 0.75] 'm{k % 4}\n")
        f.write("\t\tpoint-light [0 30 -20] [4000 4000 4000]\n")
        f.write("\t\tdiffuse-area-light [ box [0 20 0] [4 0.1 4] ] [30 30 30]\n\t]\n]\n")
    out = os.path.join(HERE, "lattice.pluscene")
    res = subprocess.run([REF_BIN, "dump-scene", urn, "0", "0", "0", out], capture_output=True, text=True)
    if res.returncode != 0:
        raise SystemExit(res.stderr)
    print(out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
