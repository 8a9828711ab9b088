#!/usr/bin/env python
"""Development tool: print a kernel's SASS around its HMMA block with the decoded scheduling fields
(stall count, write barrier, wait mask) -- `python tools/sass_loop.py <object> <mangled-name>`."""
import re, subprocess, sys
obj, fun = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", "-fun", fun, obj], capture_output=True, text=True).stdout
lines = txt.split("\n")
ins = []
i = 0
while i < len(lines):
    m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/", lines[i])
    if m and i + 1 < len(lines):
        m2 = re.search(r"/\* (0x[0-9a-f]{16}) \*/", lines[i + 1])
        if m2:
            hi = int(m2.group(1), 16)
            ins.append((m.group(1), m.group(2).strip(), (hi >> 41) & 0xF, (hi >> 46) & 7, (hi >> 52) & 0x3F))
            i += 2
            continue
    i += 1
idx = [k for k, x in enumerate(ins) if "HMMA" in x[1]]
a, b = max(0, idx[0] - int(sys.argv[3]) if len(sys.argv) > 3 else idx[0] - 60), idx[-1] + 20
tot = 0
for k in range(a, b):
    ad, t, st, wb, wt = ins[k]
    tot += st
    print(ad, t[:62].ljust(62), "st", st, "wb", wb if wb != 7 else "-", "wait", format(wt, "06b") if wt else "")
print("sum stall", tot, "n", b - a, "hmma", len(idx))
