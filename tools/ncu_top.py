#!/usr/bin/env python
"""Top sampled SASS instructions per kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda`.
  python tools/ncu_top.py src.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0

top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
i = 0
seen = set()
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Function Name":
        fn = r[1]
        hdr = rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and rows[j] and rows[j][0] not in ("File Path", "Function Name"):
            body.append(rows[j])
            j += 1
        i = j
        if "Address" not in hdr or "# Samples" not in hdr:
            continue
        ia, isamp = hdr.index("Address"), hdr.index("# Samples")
        isrc = ia + 1
        sass = [b for b in body if len(b) > isamp and b[ia].strip()]
        key = (fn, len(sass))
        if not sass or key in seen:
            continue
        seen.add(key)
        tot = sum(num(b[isamp]) for b in sass)
        print("==", fn[:150], "samples", tot)
        for b in sorted(sass, key=lambda b: -num(b[isamp]))[:top]:
            print(f"{num(b[isamp]):7d} {100.0 * num(b[isamp]) / max(tot, 1):5.1f}%  L{b[0]:>4} {b[isrc][:110]}")
    else:
        i += 1
