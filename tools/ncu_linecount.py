#!/usr/bin/env python
"""Executions per warp of given source lines (max over the SASS rows a line owns at any inline depth).
usage: ncu_linecount.py <report.ncu-rep> <library.so> <kernel substr> <source basename> <divisor> line [line ...]
With divisor = number of quads the launch decoded, the result is "times per quad"."""
import collections, csv, glob, os, re, subprocess, sys, tempfile
rep, so, kname, base, div = sys.argv[1:6]
lines = [int(a) for a in sys.argv[6:]]
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
keys = None
for cub in glob.glob(tmp + "/*.cubin"):
    txt = subprocess.run(["nvdisasm", "-gi", "-c", cub], capture_output=True, text=True).stdout
    if kname not in txt:
        continue
    infn, keys, chain, fresh = False, [], [], True
    for ln in txt.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            infn = kname in ln
            continue
        if not infn:
            continue
        mm = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if mm:
            if fresh:
                chain, fresh = [], False
            chain.append((os.path.basename(mm.group(1)), int(mm.group(2))))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            fresh = True
            keys.append([l for f, l in chain if f == base])
    break
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
ci = rows[hi].index("Instructions Executed")
body = [r for r in rows[hi + 1:] if r and r[0].startswith("0x")]
mx = collections.defaultdict(int)
for r, k in zip(body, keys):
    for l in k:
        mx[l] = max(mx[l], int(r[ci]))
for l in lines:
    print(l, round(mx[l] / float(div), 2))
