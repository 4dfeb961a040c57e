#!/usr/bin/env python
"""Device-resident decode rate of one configuration: tools/quick_rate.py dec n k crc L [frames] [snr] [f32]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import polarcodes_b200 as pb
dec, n, k, crc, L = (int(v) for v in sys.argv[1:6])
B = int(sys.argv[6]) if len(sys.argv) > 6 else 1 << 17
snr = float(sys.argv[7]) if len(sys.argv) > 7 else 2.0
f32 = len(sys.argv) > 8 and sys.argv[8] == "f32"
code = pb.Code(n, k, crc, snr)
d = pb.Decoder(n, k, crc, L, dec, pb.FP32 if (f32 and dec == 2) else pb.STRICT)
d.set_code(code.mask, code.sigma)
y = torch.empty((B, n), dtype=torch.float32, device="cuda")
tx = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
d.generate_dev(snr, B, y.data_ptr(), tx.data_ptr())
if not f32:
    y = y.double()
out = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
for _ in range(2):
    d.decode_dev(y.data_ptr(), f32, out.data_ptr(), B)
ms = [d.decode_dev(y.data_ptr(), f32, out.data_ptr(), B) for _ in range(5)]
c = d.count_errors_dev(out.data_ptr(), tx.data_ptr(), B)
ops, _ = pb.lane_ops_per_frame(code.mask, L, dec)
fps = B / (np.mean(ms) * 1e-3)
print(f"dec={dec} n={n} k={k} crc={crc} L={L} B={B} {'f32' if f32 else 'f64'}: {fps / 1e6:.3f} Mframes/s  {np.mean(ms):.3f} ms  alu_frac={fps * ops / d.alu_peak():.4f}  FER={c['frame_errors'] / B:.5f}  launch={d.info()}")
