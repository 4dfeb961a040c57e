#!/usr/bin/env python
"""Rate of the on-device frame generator alone (Philox + CRC + encode + BPSK/AWGN): tools/gen_rate.py dec n k crc L"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb
dec, n, k, crc, L = (int(v) for v in sys.argv[1:6])
B = 1 << 20 if n <= 2048 else 1 << 15
code = pb.Code(n, k, crc, 2.0)
d = pb.Decoder(n, k, crc, L, dec)
d.set_code(code.mask, code.sigma)
y = torch.empty((B, n), dtype=torch.float32, device="cuda")
tx = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
for _ in range(2):
    d.generate_dev(2.0, B, y.data_ptr(), tx.data_ptr())
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(5):
    d.generate_dev(2.0, B, y.data_ptr(), tx.data_ptr(), first_frame_id=i * B)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print(f"generate dec={dec} n={n} k={k} crc={crc}: {B / dt / 1e6:.1f} Mframes/s, {B * n * 4 / dt / 1e9:.0f} GB/s of channel values")
