#!/usr/bin/env python
"""SASS bytes per function of an object file / cubin: tools/sass_size.py FILE [filter]"""
import re, subprocess, sys
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else ""
name, n = None, {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1); n[name] = 0
    elif name and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        n[name] += 1
for k, v in n.items():
    if flt in k:
        print(v * 16, k)
