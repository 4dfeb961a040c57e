#!/usr/bin/env python
"""Device-resident throughput of the BASELINE.json configs A-E, float channel
values in HBM, packed output.  Config C is the bench.py headline.  PCB_SWEEP_ONLY=<letters> selects
configs, PCB_SWEEP_SCALE=<x> scales the batch sizes, PCB_FORCE_GENERIC=1 runs the generic kernel."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

CONFIGS = [
    ("A  SCMinsum  N=1024  K=512 (fp64 strict)", 1024, 512, 0, 1, pb.SCMINSUM, pb.STRICT, 1 << 18),
    ("A' SCMinsum  N=1024  K=512 (fp32)", 1024, 512, 0, 1, pb.SCMINSUM, 1, 1 << 18),
    ("B  SC exact  N=256   K=128", 256, 128, 0, 1, pb.SC, pb.STRICT, 1 << 18),
    ("C  SCL-8     N=1024  K=512 crc=0", 1024, 512, 0, 8, pb.SCLMINSUM, pb.STRICT, 1 << 18),
    ("C' SCL-8     N=1024  K=512 crc=16", 1024, 512, 16, 8, pb.SCLMINSUM, pb.STRICT, 1 << 18),
    ("D  SCL-32    N=2048  K=1024 crc=0", 2048, 1024, 0, 32, pb.SCLMINSUM, pb.STRICT, 1 << 16),
    ("E  SCMinsum  N=32768 K=16384 (fp64 strict)", 32768, 16384, 0, 1, pb.SCMINSUM, pb.STRICT, 1 << 16),
]
SCALE = float(os.environ.get("PCB_SWEEP_SCALE", "1"))   # batch-size multiplier
ONLY = os.environ.get("PCB_SWEEP_ONLY")                # first letters of the configs to run
for name, N, K, crc, L, dec, prec, B in CONFIGS:
    if ONLY and name[0] not in ONLY:
        continue
    B = max(32, int(B * SCALE))
    code = pb.Code(N, K, crc, 2.0)
    d = pb.Decoder(N, K, crc, L, dec, prec, device=0)
    d.set_code(code.mask, code.sigma)
    y = torch.empty((B, N), dtype=torch.float32, device="cuda")
    tx = torch.empty((B, (N + 31) // 32), dtype=torch.int32, device="cuda")
    out = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
    d.generate_dev(2.0, B, y.data_ptr(), tx.data_ptr(), seed=1, first_frame_id=0)
    ms = [d.decode_dev(y.data_ptr(), True, out.data_ptr(), B) for _ in range(3)]
    t = min(ms) * 1e-3
    print(f"{name:46s} {B / t / 1e6:9.3f} Mframes/s {B * K / t / 1e9:8.2f} info Gbit/s", flush=True)
    del d, y, tx, out
