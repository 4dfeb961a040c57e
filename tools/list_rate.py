#!/usr/bin/env python
"""List decoder (decode_scl8q.cu) frames/s over list sizes and code lengths on a device-resident float batch
(PCB_LIST_ONLY=<substring of the row name> filters).  Companion of tools/sc_rate.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

ROWS = [("C   SCL-8 1024/512", 1024, 512, 0, 8, 1 << 19), ("C16 SCL-8 1024/512 crc16", 1024, 512, 16, 8, 1 << 19),
        ("conf SCL-4 2048/1024 crc16", 2048, 1024, 16, 4, 1 << 18), ("SCL-2 1024/512", 1024, 512, 0, 2, 1 << 19),
        ("SCL-4 1024/512 crc16", 1024, 512, 16, 4, 1 << 19), ("SCL-16 1024/512 crc16", 1024, 512, 16, 16, 1 << 17),
        ("SCL-8 2048/1024 crc16", 2048, 1024, 16, 8, 1 << 18), ("SCL-8 512/256", 512, 256, 0, 8, 1 << 19),
        ("SCL-8 256/128 crc8", 256, 128, 8, 8, 1 << 19), ("SCL-16 2048/1024 crc16", 2048, 1024, 16, 16, 1 << 16),
        ("D   SCL-32 2048/1024", 2048, 1024, 0, 32, 1 << 16)]
ONLY = os.environ.get("PCB_LIST_ONLY", "")
for name, N, K, crc, L, B in ROWS:
    if ONLY and ONLY not in name:
        continue
    code = pb.Code(N, K, crc, 2.0)
    d = pb.Decoder(N, K, crc, L, pb.SCLMINSUM, pb.STRICT, device=0)
    d.set_code(code.mask, code.sigma)
    y = torch.empty((B, N), dtype=torch.float32, device="cuda")
    d.generate_dev(2.0, B, y.data_ptr(), None, seed=1)
    out = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
    ms = [d.decode_dev(y.data_ptr(), True, out.data_ptr(), B) for _ in range(4)]
    fps = B / min(ms) / 1e3
    ops, _ = pb.lane_ops_per_frame(code.mask, L, pb.SCLMINSUM)
    i = d.info()
    print(f"{name:28s} {fps:8.2f} M frames/s  alu frac {fps * 1e6 * ops / d.alu_peak():.3f}  "
          f"[warps/block {i['warps_per_block']} blocks {i['blocks']} smem {i['smem_per_block']}]", flush=True)
    d.close()
