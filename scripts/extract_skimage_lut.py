#!/usr/bin/env python
"""Write tmat_b200/csrc/skeleton_lut.inc from a scikit-image SOURCE tree (not available in this repository's
build container nor on its GPU box -- run where scikit-image 0.22's sources are at hand):

    python scripts/extract_skimage_lut.py /path/to/skimage/morphology/_skeletonize_cy.pyx [--write]

Finds the 256-entry ``lut`` initialiser of ``_fast_skeletonize`` and prints how it differs from the table shipped
here (derived from the Zhang-Suen conditions).  With --write the data part of skeleton_lut.inc is replaced; rebuild
(python -m tmat_b200.build), regenerate tests/golden (python tests/golden/make_golden.py) and re-run the tests:
kernel, oracle and fixtures all read this one file.  The neighbour-code bit order (1 NW, 2 N, 4 NE, 8 E, 16 SE,
32 S, 64 SW, 128 W) must match the correlation mask in the .pyx; the script prints the mask it finds."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "tmat_b200", "csrc", "skeleton_lut.inc")


def main():
    src = open(sys.argv[1]).read()
    m = re.search(r"lut\s*=\s*\\?\s*\[([^\]]+)\]", src, re.S)
    if not m:
        sys.exit("no `lut = [...]` initialiser found")
    vals = [int(t) for t in re.findall(r"\d+", m.group(1))]
    if len(vals) != 256 or max(vals) > 3:
        sys.exit(f"expected 256 entries in 0..3, found {len(vals)}")
    for line in src.splitlines():
        if re.search(r"\b(2|4|8|16|32|64|128)\s*\*\s*skeleton\[", line):
            print("mask:", line.strip())
    head, body = open(INC).read().split("*/", 1)
    cur = [int(t) for t in body.replace(",", " ").split()]
    diff = [i for i in range(256) if cur[i] != vals[i]]
    print(f"{len(diff)} entries differ from the shipped table:", [(i, cur[i], vals[i]) for i in diff])
    if "--write" in sys.argv:
        rows = [", ".join(str(v) for v in vals[i:i + 16]) for i in range(0, 256, 16)]
        open(INC, "w").write(head + "*/\n" + ",\n".join(rows) + "\n")
        print("wrote", INC)


if __name__ == "__main__":
    main()
