#!/usr/bin/env python
"""Aggregate an ncu source-page CSV per inlined-function chain.

usage: ncu_funcs.py <report.ncu-rep> <library.so> <kernel-mangled-substring> <source.cu> [depth] [per_frame_divisor]
Joins the SASS rows of `ncu --page source --csv` (address order) with the inline line table of
`nvdisasm -gi`, maps every (file, line) of the chain to the function of <source.cu> that contains
it and prints instructions / stall samples / top stall reasons per chain (outer>inner, at most
`depth` names).  Needs the .so compiled with -lineinfo.
"""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile


def function_ranges(path):
    """(start_line, name) for every function-like definition of the file, by a crude scan."""
    out = []
    rx = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:static\s+)?(?:__global__|__device__|__host__)[^;{]*?\b(\w+)\s*\([^;]*$")
    for i, ln in enumerate(open(path), 1):
        m = rx.match(ln)
        if m and "__launch_bounds__" not in m.group(1):
            out.append((i, m.group(1)))
        elif "__global__" in ln:
            mm = re.search(r"(\w+)\s*\(", ln.split("__launch_bounds__")[-1].split(")", 1)[-1])
            if mm:
                out.append((i, mm.group(1)))
    return out


def main():
    rep, so, kname, src = sys.argv[1:5]
    depth = int(sys.argv[5]) if len(sys.argv) > 5 else 3
    div = float(sys.argv[6]) if len(sys.argv) > 6 else 0
    fr = function_ranges(src)
    base = os.path.basename(src)

    def fn_of(file, line):
        if os.path.basename(file) != base:
            return os.path.basename(file).split(".")[0]
        name = "?"
        for s, n in fr:
            if s <= line:
                name = n
            else:
                break
        return name

    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
    keys = None
    for cub in glob.glob(os.path.join(tmp, "*.cubin")):
        txt = subprocess.run(["nvdisasm", "-gi", "-c", cub], capture_output=True, text=True).stdout
        if kname not in txt:
            continue
        infn, keys, chain = False, [], []
        fresh = True
        for ln in txt.splitlines():
            if ln.startswith("//---") and ".text." in ln:
                infn = kname in ln
                continue
            if not infn:
                continue
            mm = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if mm:
                if fresh:
                    chain = []
                    fresh = False
                chain.append((mm.group(1), int(mm.group(2))))
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                fresh = True
                names = []
                for f, l in reversed(chain):          # outermost first
                    n = fn_of(f, l)
                    if not names or names[-1] != n:
                        names.append(n)
                keys.append(">".join(names[:depth]) if names else "?")
        break
    if keys is None:
        sys.exit("kernel not found in " + so)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
    stall_cols = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    body = [r for r in rows[hdr_i + 1:] if r and r[0].startswith("0x")]
    if len(body) != len(keys):
        print(f"warning: {len(body)} profiled SASS rows vs {len(keys)} disassembled", file=sys.stderr)
    agg = {}
    tot_i = tot_s = 0
    for r, k in zip(body, keys):
        a = agg.setdefault(k, [0, 0, 0, {}])
        a[0] += int(r[ci]); a[1] += int(r[cs]); a[2] += 1
        for i, nm in stall_cols:
            v = int(r[i] or 0)
            if v:
                a[3][nm] = a[3].get(nm, 0) + v
        tot_i += int(r[ci]); tot_s += int(r[cs])
    print(f"total instructions executed {tot_i}, samples {tot_s}, SASS rows {len(body)}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        st = " ".join(f"{n}:{100 * v / max(a[1], 1):.0f}" for n, v in sorted(a[3].items(), key=lambda kv: -kv[1])[:4])
        pf = f" {a[0] / div:8.0f}/frame" if div else ""
        print(f"{k:44s} inst {100 * a[0] / tot_i:5.1f}%{pf}  samples {100 * a[1] / max(tot_s, 1):5.1f}%  sass {a[2]:5d}  {st}")


if __name__ == "__main__":
    main()
