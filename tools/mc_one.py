import sys, os
sys.path.insert(0, "/root/repo")
import polarcodes_b200 as pb
code = pb.Code(1024, 512, 16, 2.0)
d = pb.Decoder(1024, 512, 16, 8, pb.SCLMINSUM)
d.set_code(code.mask, code.sigma)
r = d.montecarlo(2.0, 1 << 18, seed=2)
print(r)
