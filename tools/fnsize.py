#!/usr/bin/env python
"""SASS rows per function label inside one kernel's text section: tools/fnsize.py <object or .so> <kernel substring>"""
import re, subprocess, sys, tempfile, os, glob
obj, key = sys.argv[1], sys.argv[2]
d = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
for cubin in glob.glob(d + "/*.cubin"):
    out = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout
    sec, lab, counts, order = None, None, {}, []
    for ln in out.splitlines():
        m = re.match(r"^//-+ \.text\.(\S+)", ln)
        if m:
            sec = m.group(1); lab = sec; continue
        if sec is None or key not in sec:
            continue
        m = re.match(r"^(\$?[_A-Za-z][^\s:]*):$", ln)
        if m and not m.group(1).startswith(".L"):
            lab = m.group(1).split("$")[-1]
        if re.match(r"^\s+/\*[0-9a-f]{4,}\*/", ln):
            k = (sec, lab)
            if k not in counts:
                counts[k] = 0; order.append(k)
            counts[k] += 1
    tot = {}
    for (s_, l_) in order:
        tot[s_] = tot.get(s_, 0) + counts[(s_, l_)]
    for s_ in tot:
        print(f"{s_[:90]}: {tot[s_]} rows = {tot[s_] * 16 / 1024:.1f} KB")
        for (a, l_) in order:
            if a == s_:
                print(f"    {counts[(a, l_)]:6d}  {l_[:100]}")
