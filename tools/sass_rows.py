#!/usr/bin/env python
"""Static SASS row count per inlined-function chain of one kernel (code size = instruction-cache
traffic for this decoder). usage: sass_rows.py <nvdisasm -gi output of the kernel> <source.cu> [depth]"""
import collections
import re
import sys

sass, srcp = sys.argv[1:3]
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 3
src = open(srcp).read().splitlines()
fr = []
rx = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:static\s+)?(?:__global__|__device__|__host__)[^;{]*?\b(\w+)\s*\([^;]*$")
for i, l in enumerate(src, 1):
    m = rx.match(l)
    if m and "__launch_bounds__" not in m.group(1):
        fr.append((i, m.group(1)))
    elif "__global__" in l:
        fr.append((i, "kernel"))


def fn(line):
    n = "?"
    for s, nm in fr:
        if s <= line:
            n = nm
        else:
            break
    return n


base = srcp.split("/")[-1]
cnt = collections.Counter()
cur, fresh = [], True
for ln in open(sass):
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        if fresh:
            cur, fresh = [], False
        cur.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        fresh = True
        names = []
        for f, l in reversed(cur):
            n = fn(l) if f == base else f.split(".")[0]
            if not names or names[-1] != n:
                names.append(n)
        cnt[">".join(names[:depth])] += 1
for k, v in cnt.most_common(30):
    print(v, k)
print(sum(cnt.values()), "rows,", sum(cnt.values()) * 16, "bytes")
