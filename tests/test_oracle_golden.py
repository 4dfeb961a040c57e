"""The C oracle (oracle/polar_oracle.c) against the committed golden vectors, which are
outputs of the unmodified reference (oracle/gen_golden.py).  CPU only."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN

DECODE = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "sc*.npz")))


def test_golden_present():
    assert len(DECODE) >= 12


@pytest.mark.parametrize("name", DECODE)
def test_decode_matches_reference_output(oracle, name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    dec, n, k, crc, L = (int(g[x]) for x in ("decoder", "n", "k", "crc", "L"))
    sol = oracle.decode(dec, n, L, k, crc, float(g["sigma"]), g["y"], g["mask"])
    assert sol.shape == g["sol"].shape
    bad = np.nonzero((sol != g["sol"]).any(axis=1))[0]
    assert bad.size == 0, f"{name}: frames {bad.tolist()} differ from the reference"


def test_encode_golden(oracle):
    g = np.load(os.path.join(GOLDEN, "encode.npz"))
    for n in (2, 4, 8, 32, 64, 256, 1024, 2048):
        u = g[f"u{n}"]
        assert (oracle.encode(u, 0) == g[f"x{n}_nat"]).all()
        assert (oracle.encode(u, 1) == g[f"x{n}_shuf"]).all()


def test_crc_known_answers(oracle):
    g = np.load(os.path.join(GOLDEN, "crc.npz"))
    # check words quoted in the reference's comments, Resources.cpp:25 and :69
    assert oracle.crc8(g["check_bits"]) == 0xF7 == int(g["check_crc8"])
    assert oracle.crc16(g["check_bits"]) == 0x29B1 == int(g["check_crc16"])
    for ln in (8, 64, 128, 512, 1024):
        for row, c8, c16 in zip(g[f"bits{ln}"], g[f"crc8_{ln}"], g[f"crc16_{ln}"]):
            assert oracle.crc8(row) == c8 and oracle.crc16(row) == c16


def test_construction_golden(oracle):
    g = np.load(os.path.join(GOLDEN, "construction.npz"))
    for key in g.files:
        parts = key.split("_")
        if parts[0] == "awgn":
            n, k, crc, snr = int(parts[1]), int(parts[2]), int(parts[3]), float(parts[4])
            got = oracle.inf_indexes_awgn(n, k + crc, oracle.sigma(n, k, snr))
        else:
            n, k = int(parts[1]), int(parts[2])
            got = oracle.inf_indexes_bec(n, k)
        assert (got == g[key]).all(), key


def test_frame_generation_consistency(oracle):
    """golden xinf/x/cw are self-consistent under the oracle's encoder, CRC and frozen-bit scatter."""
    for name in ("scl8_1024_512_crc16", "scl4_256_128_crc8", "scms_1024_512"):
        g = np.load(os.path.join(GOLDEN, name + ".npz"))
        dec, n, k, crc = (int(g[x]) for x in ("decoder", "n", "k", "crc"))
        for f in range(int(g["gen_frames"])):
            xinf = g["xinf"][f]
            assert (oracle.append_crc(xinf, k, crc) == xinf).all()
            x = oracle.add_frozen_bits(xinf, n, g["inf"])
            assert (x == g["x"][f]).all()
            assert (oracle.encode(x, dec != 3) == g["cw"][f]).all()
