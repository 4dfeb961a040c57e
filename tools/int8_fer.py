#!/usr/bin/env python
"""FER of the 8-bit min-sum mode (PCB_PREC_INT8, scales 2 << q for q = 1, 2, 3) beside the reference-exact double decoder
and the fp32 mode: SCDecoderMinsum, (1024, 512), AWGN construction at each SNR, the same on-device Philox frames for every
decoder (pcb_montecarlo).  Also frames/s of each mode (decode kernels only).  Output: the table committed as
profiles/r2_int8_fer.txt."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import polarcodes_b200 as pb

N, K = 1024, 512
FRAMES = int(os.environ.get("PCB_FER_FRAMES", 1 << 21))
modes = [("fp64 (reference arithmetic)", pb.STRICT), ("fp32", pb.FP32), ("int8 q=1 (scale 4)", pb.int8_q(1)),
         ("int8 q=2 (scale 8, conf.txt)", pb.int8_q(2)), ("int8 q=3 (scale 16)", pb.int8_q(3))]
print(f"SCDecoderMinsum N={N} K={K}, AWGN construction at the operating SNR, {FRAMES} frames per point, seed 0x5EED0000 + SNR index")
print(f"{'SNRb dB':>8s} " + " ".join(f"{m[0]:>30s}" for m in modes))
rates = {m[0]: [] for m in modes}
for si, snr in enumerate((1.0, 1.5, 2.0, 2.5, 3.0, 3.5)):
    code = pb.Code(N, K, 0, snr)
    row = []
    for name, prec in modes:
        d = pb.Decoder(N, K, 0, 1, pb.SCMINSUM, prec, device=0)
        d.set_code(code.mask, code.sigma)
        r = d.montecarlo(snr, FRAMES, seed=0x5EED0000 + si)
        row.append(f"{r['frame_errors'] / r['frames']:.3e} ({r['frame_errors']})")
        rates[name].append(r["frames"] / r["seconds"] / 1e6)
        d.close()
    print(f"{snr:8.1f} " + " ".join(f"{c:>30s}" for c in row), flush=True)
print("decode kernels only, M frames/s (float channel values from the on-device generator): " +
      ", ".join(f"{k}: {sum(v) / len(v):.1f}" for k, v in rates.items()))
