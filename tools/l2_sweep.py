#!/usr/bin/env python
"""One list-8 decode of 131072 frames (N=1024, K=512) per resident-quads-per-SM setting, in one
process, so that a single `ncu -k regex:decode_scl8q --metrics dram__bytes_read.sum,...` run
shows how the L2 behaviour of the scratch changes with the footprint.
usage: l2_sweep.py 4 8 12 16"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

N, K, B = 1024, 512, 131072
code = pb.Code(N, K, 0, 2.0)
y32 = torch.empty((B, N), dtype=torch.float32, device="cuda")
gen = pb.Decoder(N, K, 0, 8, pb.SCLMINSUM, device=0)
gen.set_code(code.mask, code.sigma)
tx = torch.empty((B, gen.words), dtype=torch.int32, device="cuda")
out = torch.empty((B, gen.words), dtype=torch.int32, device="cuda")
gen.generate_dev(2.0, B, y32.data_ptr(), tx.data_ptr(), seed=1, first_frame_id=0)
for bps in [int(a) for a in sys.argv[1:]]:
    os.environ["PCB_Q_BPS"] = str(bps)
    dec = pb.Decoder(N, K, 0, 8, pb.SCLMINSUM, device=0)
    dec.set_code(code.mask, code.sigma)
    ms = [dec.decode_dev(y32.data_ptr(), True, out.data_ptr(), B) for _ in range(3)]
    print(bps, "quads/SM:", round(B / min(ms) / 1e3, 2), "Mframes/s", flush=True)
    del dec
