#!/usr/bin/env python
"""SC decoders (configs A, A', B and other sizes): the tile-per-block kernel (decode_sct.cu) against the lane-per-frame
kernel (PCB_SC_LANE=1) on the same device batch: frames/s of both and the number of frames whose decisions differ.
Channel values are doubles for the fp64 decoders (the reference's type), floats for the fp32 mode."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

CONFIGS = [
    ("A  SCMinsum 1024/512 fp64", 1024, 512, pb.SCMINSUM, pb.STRICT, 1 << 18),
    ("A' SCMinsum 1024/512 fp32", 1024, 512, pb.SCMINSUM, pb.FP32, 1 << 18),
    ("A8 SCMinsum 1024/512 int8 q2", 1024, 512, pb.SCMINSUM, pb.int8_q(2), 1 << 18),
    ("8  SCMinsum 256/128 int8 q2", 256, 128, pb.SCMINSUM, pb.int8_q(2), 1 << 18),
    ("8  SCMinsum 2048/1024 int8 q2", 2048, 1024, pb.SCMINSUM, pb.int8_q(2), 1 << 17),
    ("B  SC exact 256/128", 256, 128, pb.SC, pb.STRICT, 1 << 18),
    ("   SCMinsum 256/128 fp64", 256, 128, pb.SCMINSUM, pb.STRICT, 1 << 18),
    ("   SCMinsum 512/256 fp64", 512, 256, pb.SCMINSUM, pb.STRICT, 1 << 18),
    ("   SCMinsum 2048/1024 fp32", 2048, 1024, pb.SCMINSUM, pb.FP32, 1 << 17),
    ("   SCMinsum 128/64 fp64", 128, 64, pb.SCMINSUM, pb.STRICT, 1 << 18),
    ("   SCMinsum 2048/1024 fp64", 2048, 1024, pb.SCMINSUM, pb.STRICT, 1 << 17),
    ("   SCMinsum 4096/2048 fp64", 4096, 2048, pb.SCMINSUM, pb.STRICT, 1 << 16),
    ("   SCMinsum 8192/4096 fp64", 8192, 4096, pb.SCMINSUM, pb.STRICT, 1 << 16),
    ("   SCMinsum 32768/16384 fp32", 32768, 16384, pb.SCMINSUM, pb.FP32, 1 << 15),
    ("E  SCMinsum 32768/16384 fp64", 32768, 16384, pb.SCMINSUM, pb.STRICT, 1 << 16),
]
ONLY = os.environ.get("PCB_SWEEP_ONLY")
SCALE = float(os.environ.get("PCB_SWEEP_SCALE", "1"))
for name, N, K, dec, prec, B in CONFIGS:
    if ONLY and name[0] not in ONLY:
        continue
    B = max(32, int(B * SCALE)) + 5            # a ragged last tile
    code = pb.Code(N, K, 0, 2.0)
    res = {}
    int8 = prec not in (pb.STRICT, pb.FP32)
    for variant in ("tile",) if int8 else ("tile", "lane"):
        if variant == "lane":
            os.environ["PCB_SC_LANE"] = "1"
        try:
            d = pb.Decoder(N, K, 0, 1, dec, prec, device=0)
        finally:
            os.environ.pop("PCB_SC_LANE", None)
        d.set_code(code.mask, code.sigma)
        y32 = torch.empty((B, N), dtype=torch.float32, device="cuda")
        d.generate_dev(2.0, B, y32.data_ptr(), None, seed=1, first_frame_id=0)
        f32 = prec == pb.FP32
        y = y32 if f32 else y32.double()
        out = torch.zeros((B, d.words), dtype=torch.int32, device="cuda")
        ms = [d.decode_dev(y.data_ptr(), f32, out.data_ptr(), B) for _ in range(4)]
        res[variant] = (B / (min(ms) * 1e-3) / 1e6, out.clone(), d.info())
        del d
    if int8:
        res["lane"] = (0.0, res["tile"][1], None)
    diff = (res["tile"][1] != res["lane"][1]).any(dim=1).sum().item()
    i = res["tile"][2]
    print(f"{name:30s} tile {res['tile'][0]:8.2f}  lane {res['lane'][0]:8.2f} Mframes/s   frames that differ: {diff} of {B}   "
          f"[W {i['warps_per_block']} vs {i['first_smem_layer']} blocks {i['blocks']} smem {i['smem_per_block']}]", flush=True)
