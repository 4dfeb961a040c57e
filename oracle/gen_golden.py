"""TEST INFRASTRUCTURE — generates tests/golden/*.npz from the UNMODIFIED reference
(oracle/_ref/libpolar_ref.so, built by oracle/Makefile from /root/reference).

Run in the build container (where /root/reference exists):  python oracle/gen_golden.py
The reference ships no golden vectors of its own, so these are outputs of the
reference itself on seeded inputs: eng.seed(1234) (Resources.cpp:15 is a plain global), frames
drawn by the reference's own loop body (PolarCodes.cpp:133-158).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.pyoracle import Ref  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# name, decoder, n, k, crc, L, snr_db, zcomputer, frames
DECODE_CASES = [
    ("scms_1024_512", 2, 1024, 512, 0, 1, 2.0, 2, 16),
    ("scms_1024_512_bec", 2, 1024, 512, 0, 1, 1.5, 1, 8),
    ("sc_256_128", 1, 256, 128, 0, 1, 2.0, 2, 32),
    ("scl8_1024_512_crc0", 3, 1024, 512, 0, 8, 2.0, 2, 16),
    ("scl8_1024_512_crc16", 3, 1024, 512, 16, 8, 1.5, 2, 16),
    ("scl32_2048_1024_crc16", 3, 2048, 1024, 16, 32, 2.0, 2, 4),
    ("scl32_2048_1024_crc0", 3, 2048, 1024, 0, 32, 2.0, 2, 2),
    ("scl4_256_128_crc8", 3, 256, 128, 8, 4, 1.5, 2, 32),
    ("scl2_64_32_crc0", 3, 64, 32, 0, 2, 1.0, 2, 64),
    ("scl16_128_64_crc0", 3, 128, 64, 0, 16, 2.0, 2, 32),
    ("scl1_128_64_crc8", 3, 128, 64, 8, 1, 2.0, 2, 32),
    ("scms_32768_16384", 2, 32768, 16384, 0, 1, 2.0, 2, 1),
]


def edge_frames(n, rng):
    """Targeted inputs: all-zero, +-0 mix, exact ties (small integer grid), saturated, tiny."""
    e = []
    e.append(np.zeros(n))
    z = np.zeros(n)
    z[::2] = -0.0
    e.append(z)
    e.append(rng.integers(-2, 3, n).astype(np.float64))          # many exact ties and zeros
    e.append(rng.integers(-1, 2, n).astype(np.float64) * 0.5)
    e.append(np.where(rng.random(n) < 0.5, 1e30, -1e30))          # saturated magnitudes
    e.append(rng.standard_normal(n) * 1e-30)                      # products underflow in fp32
    e.append(np.where(rng.random(n) < 0.5, 3.0e38, -3.0e38))      # metric overflows to inf in fp32
    e.append(np.ones(n))
    e.append(-np.ones(n))
    return np.stack(e)


def main():
    os.makedirs(OUT, exist_ok=True)
    r = Ref()
    rng = np.random.default_rng(20261104)
    for name, dec, n, k, crc, L, snr, zc, frames in DECODE_CASES:
        r.seed(1234)
        sigma, inf, mask = r.make_code(n, k, crc, snr, zc)
        xinf, x, cw, y = r.gen_frames(n, k, crc, inf, sigma, dec != 3, frames)
        if n <= 2048:
            y = np.concatenate([y, edge_frames(n, rng)])
        sol = r.decode(dec, n, L, k, crc, sigma, y, mask)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), decoder=dec, n=n, k=k, crc=crc, L=L, snr_db=snr,
                            zcomputer=zc, sigma=sigma, inf=inf, mask=mask, xinf=xinf, x=x, cw=cw, y=y, sol=sol,
                            gen_frames=frames)
        print(name, y.shape, sol.shape)

    # encoder + CRC + construction known answers
    enc = {}
    for n in (2, 4, 8, 32, 64, 256, 1024, 2048):
        u = rng.integers(0, 2, (6, n)).astype(np.int32)
        enc[f"u{n}"] = u
        enc[f"x{n}_nat"] = r.encode(u, 0)
        enc[f"x{n}_shuf"] = r.encode(u, 1)
    np.savez_compressed(os.path.join(OUT, "encode.npz"), **enc)

    crcd = {}
    msg = np.array([(b >> (7 - i)) & 1 for b in b"123456789" for i in range(8)], np.int32)
    crcd["check_bits"] = msg
    crcd["check_crc8"] = r.crc8(msg)          # 0xF7  (Resources.cpp:25)
    crcd["check_crc16"] = r.crc16(msg)        # 0x29B1 (Resources.cpp:69)
    for ln in (8, 64, 128, 512, 1024):
        b = rng.integers(0, 2, (8, ln)).astype(np.int32)
        crcd[f"bits{ln}"] = b
        crcd[f"crc8_{ln}"] = np.array([r.crc8(row) for row in b], np.uint32)
        crcd[f"crc16_{ln}"] = np.array([r.crc16(row) for row in b], np.uint32)
    np.savez_compressed(os.path.join(OUT, "crc.npz"), **crcd)

    cons = {}
    for n, k, crc in ((1024, 512, 0), (1024, 512, 16), (2048, 1024, 16), (256, 128, 0), (256, 128, 8),
                      (64, 32, 0), (32768, 16384, 0)):
        for snr in (1.0, 2.0, 4.0):
            s = r.sigma(n, k, snr)
            cons[f"awgn_{n}_{k}_{crc}_{snr}"] = r.inf_indexes_awgn(n, k + crc, s)
        cons[f"bec_{n}_{k}"] = r.inf_indexes_bec(n, k)
    np.savez_compressed(os.path.join(OUT, "construction.npz"), **cons)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
