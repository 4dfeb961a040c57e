"""CPU-only checks of the product's host side: the C-ABI library loads and exports every
symbol include/polar_b200.h declares, the host-side code construction / CRC / frame assembly
match the golden vectors of the reference, and compute entry points fail loudly (no CPU
fallback) when there is no CUDA device."""
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT


@pytest.fixture(scope="module")
def pb():
    import polarcodes_b200
    from polarcodes_b200 import build
    build.build()
    return polarcodes_b200


def test_header_symbols_exported(pb):
    hdr = open(os.path.join(ROOT, "include", "polar_b200.h")).read()
    declared = set(re.findall(r"\b(pcb_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(pb.api.SIGNATURES), declared ^ set(pb.api.SIGNATURES)
    L = pb.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert b"sm_100a" in L.pcb_version()


def test_no_cpu_fallback(pb):
    if pb.lib().pcb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pb.PolarError, match="no CUDA device"):
        pb.Decoder(1024, 512, 16, 8, pb.SCLMINSUM)


def test_product_never_touches_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "polarcodes_b200")):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "pyoracle" not in src and "polar_oracle" not in src, f


def test_construction_golden(pb):
    g = np.load(os.path.join(GOLDEN, "construction.npz"))
    L = pb.lib()
    for key in g.files:
        parts = key.split("_")
        if parts[0] == "awgn":
            n, k, crc, snr = int(parts[1]), int(parts[2]), int(parts[3]), float(parts[4])
            got = np.zeros(k + crc, np.int32)
            assert L.pcb_inf_indexes_awgn(n, k + crc, L.pcb_sigma(n, k, snr), got) == 0
        else:
            n, k = int(parts[1]), int(parts[2])
            got = np.zeros(k, np.int32)
            assert L.pcb_inf_indexes_bec(n, k, got) == 0
        assert (got == g[key]).all(), key


def test_code_matches_golden_masks(pb):
    for name in ("scl8_1024_512_crc16", "scms_1024_512_bec", "scl32_2048_1024_crc0", "sc_256_128"):
        g = np.load(os.path.join(GOLDEN, name + ".npz"))
        c = pb.Code(int(g["n"]), int(g["k"]), int(g["crc"]), float(g["snr_db"]), int(g["zcomputer"]))
        assert c.sigma == float(g["sigma"])
        assert (c.inf == g["inf"]).all() and (c.mask == g["mask"]).all()


def test_crc_known_answers(pb):
    g = np.load(os.path.join(GOLDEN, "crc.npz"))
    assert pb.crc8(g["check_bits"]) == 0xF7      # Resources.cpp:25
    assert pb.crc16(g["check_bits"]) == 0x29B1   # Resources.cpp:69
    for ln in (8, 64, 128, 512, 1024):
        for row, c8, c16 in zip(g[f"bits{ln}"], g[f"crc8_{ln}"], g[f"crc16_{ln}"]):
            assert pb.crc8(row) == c8 and pb.crc16(row) == c16


def test_frame_assembly_golden(pb):
    for name in ("scl8_1024_512_crc16", "scl4_256_128_crc8", "scms_1024_512"):
        g = np.load(os.path.join(GOLDEN, name + ".npz"))
        n, k, crc = (int(g[x]) for x in ("n", "k", "crc"))
        for f in range(int(g["gen_frames"])):
            xinf = g["xinf"][f]
            assert (pb.append_crc(xinf, k, crc) == xinf).all()
            assert (pb.add_frozen_bits(xinf, n, g["inf"]) == g["x"][f]).all()


def test_bad_arguments(pb):
    L = pb.lib()
    out = np.zeros(4, np.int32)
    assert L.pcb_inf_indexes_awgn(100, 4, 1.0, out) < 0 and b"bad n/k" in L.pcb_last_error()
    with pytest.raises(pb.PolarError):
        pb.append_crc(np.zeros(20, np.int32), 12, 8)   # k % 8 != 0
    for args in ((0, 1000, 500, 0, 8, 3, 0), (0, 1024, 512, 4, 8, 3, 0), (0, 1024, 512, 0, 3, 3, 0),
                 (0, 1024, 512, 0, 8, 7, 0), (0, 1024, 0, 0, 8, 3, 0)):
        assert not L.pcb_create(*args)
        assert L.pcb_last_error()
