#!/usr/bin/env python
"""On-device Monte-Carlo rate (Philox BPSK/AWGN + encode + decode + error count, nothing crosses
PCIe) of the BASELINE.json configs: frames/s of pcb_montecarlo, FER and BER at 2 dB."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import polarcodes_b200 as pb

CONFIGS = [
    ("A  SCMinsum  N=1024  K=512", 1024, 512, 0, 1, pb.SCMINSUM, 1 << 21),
    ("B  SC exact  N=256   K=128", 256, 128, 0, 1, pb.SC, 1 << 21),
    ("C  SCL-8     N=1024  K=512 crc=16", 1024, 512, 16, 8, pb.SCLMINSUM, 1 << 21),
    ("D  SCL-32    N=2048  K=1024 crc=16", 2048, 1024, 16, 32, pb.SCLMINSUM, 1 << 18),
    ("E  SCMinsum  N=32768 K=16384", 32768, 16384, 0, 1, pb.SCMINSUM, 42624 * 4),
]
ONLY = os.environ.get("PCB_SWEEP_ONLY")
for name, N, K, crc, L, dec, frames in CONFIGS:
    if ONLY and name[0] not in ONLY:
        continue
    code = pb.Code(N, K, crc, 2.0)
    d = pb.Decoder(N, K, crc, L, dec, pb.STRICT, device=0)
    d.set_code(code.mask, code.sigma)
    d.montecarlo(2.0, frames // 4, seed=1)          # warm-up, buffers
    t0 = time.time()
    r = d.montecarlo(2.0, frames, seed=2)
    dt = time.time() - t0
    bits = K if dec == pb.SCLMINSUM else N
    print(f"{name:36s} {frames / dt / 1e6:8.3f} Mframes/s wall ({frames / r['seconds'] / 1e6:8.3f} decode only)  "
          f"FER {r['frame_errors'] / frames:.3e}  BER {r['bit_errors'] / frames / bits:.3e}", flush=True)
    d.close()
