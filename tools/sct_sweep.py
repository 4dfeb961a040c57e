#!/usr/bin/env python
"""Launch-shape sweep of the tile-per-block SC kernel (decode_sct.cu): warps per tile (PCB_T_W) x first shared-memory level
(PCB_T_VS) for a list of codes; prints frames/s per shape and the best one.  PCB_SWEEP_ONLY=<substring> filters codes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

CODES = [("ms64", 256, pb.SCMINSUM, pb.STRICT), ("ms64", 512, pb.SCMINSUM, pb.STRICT), ("ms64", 1024, pb.SCMINSUM, pb.STRICT),
         ("ms64", 2048, pb.SCMINSUM, pb.STRICT), ("ms32", 256, pb.SCMINSUM, pb.FP32), ("ms32", 1024, pb.SCMINSUM, pb.FP32),
         ("ms32", 2048, pb.SCMINSUM, pb.FP32), ("exact", 256, pb.SC, pb.STRICT), ("exact", 1024, pb.SC, pb.STRICT)]
ONLY = os.environ.get("PCB_SWEEP_ONLY", "")
os.environ["PCB_T_MINBPS"] = "99"
for name, N, dec, prec in CODES:
    if ONLY and ONLY not in f"{name}{N}":
        continue
    B = (1 << 28) // (N * 1024) * 1024 // (4 if name == "exact" else 1)
    code = pb.Code(N, N // 2, 0, 2.0)
    f32 = prec == pb.FP32
    res = []
    y = None
    for W in (1, 2, 4):
        for VS in range(2, N.bit_length() - 4):
            os.environ["PCB_T_W"], os.environ["PCB_T_VS"] = str(W), str(VS)
            try:
                d = pb.Decoder(N, N // 2, 0, 1, dec, prec, device=0)
            except pb.PolarError:
                continue
            d.set_code(code.mask, code.sigma)
            i = d.info()
            if i["first_smem_layer"] != VS or i["warps_per_block"] != W:
                del d
                continue
            if y is None:
                y32 = torch.empty((B, N), dtype=torch.float32, device="cuda")
                d.generate_dev(2.0, B, y32.data_ptr(), None, seed=1, first_frame_id=0)
                y = y32 if f32 else y32.double()
                out = torch.zeros((B, d.words), dtype=torch.int32, device="cuda")
            ms = min(d.decode_dev(y.data_ptr(), f32, out.data_ptr(), B) for _ in range(3))
            res.append((B / ms / 1e3, W, VS, i["blocks"] // i["sm_count"], i["smem_per_block"]))
            del d
    res.sort(reverse=True)
    print(f"{name} n={N} B={B}: " + "  ".join(f"W{w}/VS{v}:{r:.1f}({t}t)" for r, w, v, t, s in res[:6]), flush=True)
    del y, out
