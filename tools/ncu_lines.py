#!/usr/bin/env python
"""Aggregate an ncu source-page CSV (SASS rows) per CUDA source line.

usage: ncu_lines.py <report.ncu-rep> <library.so> <kernel-mangled-substring> [topN]
Joins the SASS rows of `ncu --page source --csv` (in address order) with the line table that
`nvdisasm -g` prints for the same function, and prints instructions executed / stall samples per
source line.  Needs the .so compiled with -lineinfo.
"""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, so, kname = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
    lines = None
    for cub in glob.glob(os.path.join(tmp, "*.cubin")):
        txt = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
        if kname not in txt:
            continue
        cur, infn, lines = None, False, []
        for ln in txt.splitlines():
            if ln.startswith("//---") and ".text." in ln:
                infn = kname in ln
                continue
            if not infn:
                continue
            mm = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if mm:
                cur = (os.path.basename(mm.group(1)), int(mm.group(2)))
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                lines.append(cur)
        break
    if lines is None:
        sys.exit("kernel not found in " + so)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
    body = [r for r in rows[hdr_i + 1:] if r and r[0].startswith("0x")]
    if len(body) != len(lines):
        print(f"warning: {len(body)} profiled SASS rows vs {len(lines)} disassembled", file=sys.stderr)
    agg = {}
    tot_i = tot_s = 0
    for r, l in zip(body, lines):
        a = agg.setdefault(l, [0, 0, 0])
        a[0] += int(r[ci]); a[1] += int(r[cs]); a[2] += 1
        tot_i += int(r[ci]); tot_s += int(r[cs])
    print(f"total instructions executed {tot_i}, samples {tot_s}, SASS rows {len(body)}")
    for l, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{str(l):40s} inst {a[0]:>14d} {100 * a[0] / tot_i:5.1f}%  samples {a[1]:>8d} {100 * a[1] / max(tot_s, 1):5.1f}%  sass {a[2]}")


if __name__ == "__main__":
    main()
