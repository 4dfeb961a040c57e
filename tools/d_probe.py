import os, sys
sys.path.insert(0, '/root/repo')
import torch, polarcodes_b200 as pb
for name, N, K, crc, L, B in (("D SCL-32 2048", 2048, 1024, 0, 32, 1 << 16), ("SCL-16 2048 crc16", 2048, 1024, 16, 16, 1 << 16)):
    code = pb.Code(N, K, crc, 2.0)
    d = pb.Decoder(N, K, crc, L, pb.SCLMINSUM, pb.STRICT, device=0); d.set_code(code.mask, code.sigma)
    y = torch.empty((B, N), dtype=torch.float32, device="cuda")
    d.generate_dev(2.0, B, y.data_ptr(), None, seed=1)
    out = torch.empty((B, d.words), dtype=torch.int32, device="cuda")
    ms = [d.decode_dev(y.data_ptr(), True, out.data_ptr(), B) for _ in range(4)]
    print(name, "%.3f Mframes/s" % (B / min(ms) / 1e3), d.info(), flush=True)
