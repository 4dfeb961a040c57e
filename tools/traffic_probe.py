#!/usr/bin/env python
"""DRAM traffic and executed instructions per decode launch for every bench configuration.

  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum \\
      --clock-control none --csv --log-file gpurun_out/r2_traffic.csv python tools/traffic_probe.py run
  python tools/traffic_probe.py parse gpurun_out/r2_traffic.csv profiles/r2_traffic.json

`run` decodes ONE device-resident batch per configuration (a generate_kernel launch separates the
configurations in the ncu launch list); `parse` sums the decode-side kernels of each configuration.
bench.py reads profiles/r2_traffic.json for `roofline.traffic` and `configs[*].dram_over_algorithmic`.
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# key, decoder, n, k, crc, L, fp32 precision, double input, frames  (same keys as bench.py's `configs`, "C" = headline)
CONFIGS = [
    ("C", 3, 1024, 512, 0, 8, 0, True, 1 << 18),
    ("A", 2, 1024, 512, 0, 1, 0, True, 1 << 18),
    ("A_fp32", 2, 1024, 512, 0, 1, 1, False, 1 << 18),
    ("A_int8", 2, 1024, 512, 0, 1, 2 | (2 << 8), True, 1 << 18),
    ("B", 1, 256, 128, 0, 1, 0, True, 1 << 18),
    ("C_crc16", 3, 1024, 512, 16, 8, 0, True, 1 << 18),
    ("conf_txt", 3, 2048, 1024, 16, 4, 0, True, 1 << 17),
    ("D", 3, 2048, 1024, 0, 32, 0, True, 1 << 16),
    ("E", 2, 32768, 16384, 0, 1, 0, True, 85248),
]


def run():
    import torch
    import polarcodes_b200 as pb
    for key, decoder, n, k, crc, L, prec, dbl, B in CONFIGS:
        code = pb.Code(n, k, crc, 2.0)
        dec = pb.Decoder(n, k, crc, L, decoder, prec)
        dec.set_code(code.mask, code.sigma)
        y = torch.empty((B, n), dtype=torch.float32, device="cuda")
        tx = torch.empty((B, dec.words), dtype=torch.int32, device="cuda")
        dec.generate_dev(2.0, B, y.data_ptr(), tx.data_ptr())
        if dbl:
            y = y.double()
        sol = torch.empty((B, dec.words), dtype=torch.int32, device="cuda")
        dec.decode_dev(y.data_ptr(), not dbl, sol.data_ptr(), B)
        torch.cuda.synchronize()
        dec.close()
        del y, tx, sol
        torch.cuda.empty_cache()
        print(key, "done", flush=True)


def parse(src, dst):
    rows = []
    with open(src, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    for r in csv.DictReader(lines):
        rows.append(r)
    # ncu --csv (raw page off): one row per (launch, metric)
    launches = {}
    order = []
    for r in rows:
        i = int(r["ID"])
        if i not in launches:
            launches[i] = {"name": r["Kernel Name"]}
            order.append(i)
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        name = r["Metric Name"]
        if name.startswith("dram__bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)
        if name.startswith("gpu__time"):
            v *= {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9, "second": 1, "s": 1}.get(unit, 1)
        launches[i][name] = v
    out, ci = {}, -1
    for i in order:
        L = launches[i]
        nm = L["name"]
        if "generate_kernel" in nm:
            ci += 1
            continue
        if ci < 0 or ci >= len(CONFIGS):
            continue
        if not any(t in nm for t in ("decode_", "prepass", "transpose", "expand", "scl_select", "quantise")):
            continue                                        # torch's own kernels (the float -> double widening)
        key, B = CONFIGS[ci][0], CONFIGS[ci][8]
        o = out.setdefault(key, {"frames_per_launch": B, "dram_bytes_read": 0.0, "dram_bytes_write": 0.0, "kernel_s": 0.0,
                                 "warp_instructions": 0.0, "kernels": []})
        o["dram_bytes_read"] += L.get("dram__bytes_read.sum", 0.0)
        o["dram_bytes_write"] += L.get("dram__bytes_write.sum", 0.0)
        o["kernel_s"] += L.get("gpu__time_duration.sum", 0.0)
        o["warp_instructions"] += L.get("smsp__inst_executed.sum", 0.0)
        o["kernels"].append({"name": nm.split("(")[0][:80], "ms": 1e3 * L.get("gpu__time_duration.sum", 0.0),
                             "dram_GB": (L.get("dram__bytes_read.sum", 0.0) + L.get("dram__bytes_write.sum", 0.0)) / 1e9})
    for key, o in out.items():
        o["source"] = (f"{os.path.basename(src)}: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,"
                       "smsp__inst_executed.sum --clock-control none, one launch of tools/traffic_probe.py run (cold caches, serialised)")
    json.dump(out, open(dst, "w"), indent=1)
    for key, o in out.items():
        print(key, f"{(o['dram_bytes_read'] + o['dram_bytes_write']) / o['frames_per_launch']:.0f} B/frame",
              f"{o['warp_instructions'] / o['frames_per_launch']:.0f} warp-instr/frame", f"{1e3 * o['kernel_s']:.2f} ms")


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run()
    else:
        parse(sys.argv[2], sys.argv[3])
