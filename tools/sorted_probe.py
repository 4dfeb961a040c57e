#!/usr/bin/env python
"""Does grouping frames of similar difficulty into the same warp pay?  The four frames of a warp
take the slow fork path together, so a quad of easy frames is faster than a mixed one.  Decodes the
same batch in its natural order and sorted by cheap per-frame proxies (config C, float input)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

N, K, crc, L, B = 1024, 512, int(os.environ.get("CRC", "0")), 8, 1 << 18
code = pb.Code(N, K, crc, 2.0)
d = pb.Decoder(N, K, crc, L, pb.SCLMINSUM, pb.STRICT, device=0)
d.set_code(code.mask, code.sigma)
y = torch.empty((B, N), dtype=torch.float32, device="cuda")
tx = torch.empty((B, (N + 31) // 32), dtype=torch.int32, device="cuda")
out = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
d.generate_dev(2.0, B, y.data_ptr(), tx.data_ptr(), seed=1, first_frame_id=0)


def run(name, yy):
    ms = [d.decode_dev(yy.data_ptr(), True, out.data_ptr(), B) for _ in range(4)]
    print(f"{name:40s} {B / min(ms) / 1e3:9.3f} Mframes/s", flush=True)


run("natural order", y)
a = y.abs()
info = torch.tensor((code.mask == -1) if hasattr(code.mask, "__eq__") else [m == -1 for m in code.mask], device="cuda")
proxies = {
    "sum |y|": a.sum(1),
    "count |y| < 0.3": (a < 0.3).sum(1).float(),
    "count |y| < 0.5": (a < 0.5).sum(1).float(),
    "sum (|y|-1)^2": ((a - 1) ** 2).sum(1),
    "sum min(|y|,1)": a.clamp(max=1.0).sum(1),
    "random permutation": torch.rand(B, device="cuda"),
}
for name, key in proxies.items():
    perm = torch.argsort(key)
    run("sorted by " + name, y[perm].contiguous())
