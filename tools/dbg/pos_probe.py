"""Which frames of config A (float channel values, fp64 decoder) differ between the tile kernel and the generic kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import polarcodes_b200 as pb

n, k, B = 1024, 512, int(os.environ.get('PROBE_B', 1 << 15))
code = pb.Code(n, k, 0, 2.0)
def handle(generic=False):
    if generic: os.environ["PCB_FORCE_GENERIC"] = "1"
    try: d = pb.Decoder(n, k, 0, 1, pb.SCMINSUM, pb.STRICT, device=0)
    finally: os.environ.pop("PCB_FORCE_GENERIC", None)
    d.set_code(code.mask, code.sigma)
    return d
fast, slow = handle(), handle(True)
print(fast.info())
y = torch.empty((B, n), dtype=torch.float32, device="cuda")
fast.generate_dev(2.0, B, y.data_ptr(), None, seed=77, first_frame_id=0)
def dec(d, y, f32=True):
    out = torch.full((y.shape[0], d.words), 0x5a5a5a5a, dtype=torch.int32, device="cuda")
    d.decode_dev(y.data_ptr(), f32, out.data_ptr(), y.shape[0]); torch.cuda.synchronize()
    return out
want = dec(slow, y)
for rep in range(3):
    got = dec(fast, y)
    bad = torch.nonzero((got != want).any(dim=1)).flatten()
    print("float y: frames that differ", bad.numel(), bad[:16].tolist(), "untouched rows", (got == 0x5a5a5a5a).all(dim=1).sum().item())
    if bad.numel():
        f = bad[0].item(); print(" words differing in frame", f, torch.nonzero(got[f] != want[f]).flatten()[:16].tolist())
yd = y.double()
for rep in range(3):
    got = dec(fast, yd, False)
    bad = torch.nonzero((got != want).any(dim=1)).flatten()
    print("double y: frames that differ", bad.numel(), bad[:4].tolist(), bad[-4:].tolist())
    if bad.numel():
        f = bad[0].item(); w = torch.nonzero(got[f] != want[f]).flatten(); print(" words differing in frame", f, w.numel(), w[:16].tolist())
        f = bad[-1].item(); w = torch.nonzero(got[f] != want[f]).flatten(); print(" words differing in frame", f, w.numel(), w[:16].tolist())
for rep in range(2):
    got = dec(fast, y)
    bad = torch.nonzero((got != want).any(dim=1)).flatten()
    print("float y again: frames that differ", bad.numel(), bad[:4].tolist(), bad[-4:].tolist())
