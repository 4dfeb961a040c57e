#!/usr/bin/env python
"""Throughput of the list-8 decoder (device-resident float input) for several code sizes.
usage: size_sweep.py [crc]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

crc = int(sys.argv[1]) if len(sys.argv) > 1 else 0
for N, K in ((256, 128), (512, 256), (1024, 512), (2048, 1024)):
    B = 131072 * 1024 // N
    code = pb.Code(N, K, crc, 2.0)
    dec = pb.Decoder(N, K, crc, 8, pb.SCLMINSUM, device=0)
    dec.set_code(code.mask, code.sigma)
    y = torch.empty((B, N), dtype=torch.float32, device="cuda")
    tx = torch.empty((B, dec.words), dtype=torch.int32, device="cuda")
    out = torch.empty((B, dec.words), dtype=torch.int32, device="cuda")
    dec.generate_dev(2.0, B, y.data_ptr(), tx.data_ptr(), seed=1, first_frame_id=0)
    ms = [dec.decode_dev(y.data_ptr(), True, out.data_ptr(), B) for _ in range(4)]
    info = dec.info()
    print(f"N={N} K={K} crc={crc}: {B / min(ms) / 1e3:.2f} Mframes/s, {B * K / min(ms) / 1e6:.2f} Gbit/s, blocks {info['blocks']}, smem/block {info['smem_per_block']}", flush=True)
    del dec
