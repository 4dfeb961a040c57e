#!/usr/bin/env python
"""Config E (SCDecoderMinsum N=32768, fp64): the round-1 lane-per-frame kernel against the tile kernel with its levels in the
global slice (PCB_T_BIG=1) for a few launch shapes, same batch, decisions compared."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polarcodes_b200 as pb

N, K, B = 32768, 16384, int(os.environ.get("PCB_E_FRAMES", 1 << 16))
code = pb.Code(N, K, 0, 2.0)
y32 = torch.empty((B, N), dtype=torch.float32, device="cuda")
ref = None
shapes = [None] + [(4, 9), (4, 10), (8, 9)]
for sh in shapes:
    for k in ("PCB_T_BIG", "PCB_T_W", "PCB_T_VS"):
        os.environ.pop(k, None)
    if sh:
        os.environ.update(PCB_T_BIG="1", PCB_T_W=str(sh[0]), PCB_T_VS=str(sh[1]))
    d = pb.Decoder(N, K, 0, 1, pb.SCMINSUM, pb.STRICT, device=0)
    d.set_code(code.mask, code.sigma)
    if ref is None:
        d.generate_dev(2.0, B, y32.data_ptr(), None, seed=1)
        y = y32.double()
        del y32
    out = torch.zeros((B, d.words), dtype=torch.int32, device="cuda")
    ms = [d.decode_dev(y.data_ptr(), False, out.data_ptr(), B) for _ in range(2)]
    if ref is None:
        ref = out.clone()
    diff = (out != ref).any(dim=1).sum().item()
    i = d.info()
    print(f"{'lane kernel' if not sh else 'tile W%d VS%d' % sh:14s} {B / min(ms) / 1e3:6.3f} M frames/s  differ {diff}  blocks {i['blocks']} smem {i['smem_per_block']}", flush=True)
    d.close()
