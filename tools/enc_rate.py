#!/usr/bin/env python
"""Packed encoder (pcb_encode_batch_dev) frames/s and HBM GB/s over code lengths, natural order and with ReverseShuffle."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

ONLY = os.environ.get("PCB_ENC_ONLY")
for n, B in ((256, 1 << 22), (1024, 1 << 22), (2048, 1 << 21), (4096, 1 << 20), (8192, 1 << 19), (32768, 1 << 17), (65536, 1 << 16)):
    if ONLY and str(n) != ONLY:
        continue
    d = pb.Decoder(n, n // 2, 0, 1, pb.SCMINSUM, device=0)
    u = torch.randint(-2**31, 2**31 - 1, (B, n // 32), dtype=torch.int32, device="cuda")
    x = torch.zeros_like(u)
    for sh in (0, 1):
        ms = min(d.encode_dev(u.data_ptr(), x.data_ptr(), sh, B) for _ in range(5))
        print(f"n={n} shuffle={sh}: {B / ms / 1e3:.0f} M frames/s  {2 * B * n / 8 / ms / 1e6:.0f} GB/s", flush=True)
    d.close()
