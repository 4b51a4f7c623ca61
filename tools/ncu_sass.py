#!/usr/bin/env python
"""Barrier waits, tensor / TMA issues and hot instructions of one kernel launch, from
`ncu -i X.ncu-rep --page source --csv --print-source sass > sass.csv`.
  python tools/ncu_sass.py sass.csv <launch index, 0-based> [min share %]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = int(sys.argv[2])
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
starts.append(len(rows))
a, b = starts[want], starts[want + 1]
hdr, body = rows[a + 1], rows[a + 2:b]
isamp, isrc, iex = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
tot = sum(num(r[isamp]) for r in body)
print("==", rows[a][1][:160], "| samples", tot, "| instructions", len(body))
for n, r in enumerate(body):
    s = r[isrc]
    if "TRYWAIT" in s or num(r[isamp]) > tot * thr / 100 or "UTCHMMA" in s or "UTMALDG" in s or "UTCBAR" in s:
        print(f"{n:5d} {num(r[isamp]):6d} {100 * num(r[isamp]) / max(tot, 1):5.1f}% ex={r[iex]:>9} {s[:100]}")
