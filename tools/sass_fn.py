#!/usr/bin/env python
"""Print a window of one kernel's SASS around its tensor-core issues (from `cuobjdump -sass`).
  python tools/sass_fn.py file.sass <substring of the mangled name> [before] [after]"""
import re
import sys

txt = open(sys.argv[1]).read()
pat = sys.argv[2]
before = int(sys.argv[3]) if len(sys.argv) > 3 else 40
after = int(sys.argv[4]) if len(sys.argv) > 4 else 25
for f in txt.split("Function : ")[1:]:
    name = f.split("\n")[0]
    if pat not in name:
        continue
    lines = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln) for ln in f.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", ln)]
    idx = [i for i, ln in enumerate(lines) if "UTCHMMA" in ln]
    print(name, "| instructions:", len(lines), "| UTCHMMA at", idx)
    if idx:
        for ln in lines[max(0, idx[0] - before):idx[min(3, len(idx) - 1)] + after]:
            print(ln[:110])
    break
