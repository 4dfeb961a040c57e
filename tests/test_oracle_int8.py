"""The CPU definition of the 8-bit min-sum mode (oracle/polar_oracle.c: po_q8 / po_f8 / po_g8 / po_sc8_decode_batch) against
the reference's own 8-bit sketches (SCDecoderMinsum.cpp:17-30, through oracle/_ref), and basic sanity of the decoder."""
import numpy as np
import pytest

import polarcodes_b200 as pb


def test_f8_g8_equal_the_reference_char_functions(oracle, ref):
    o, r = oracle.lib, ref.lib
    for a in range(-127, 128):
        for b in range(-127, 128):
            assert o.po_f8(a, b) == r.ref_decodef_char(a, b), (a, b)
            for c in (0, 1):
                v = b + a * (1 - 2 * c)
                if abs(v) <= 127:                                   # where Decodeg_char does not wrap the two are the same sum
                    assert o.po_g8(a, b, c) == r.ref_decodeg_char(a, b, c), (a, b, c)
                else:
                    assert o.po_g8(a, b, c) == (127 if v > 0 else -127)


def test_q8_rounds_to_even_and_saturates(oracle):
    o = oracle.lib
    o.po_q8.argtypes = [__import__("ctypes").c_double, __import__("ctypes").c_int]
    assert [o.po_q8(v, 2) for v in (0.0, 0.0625, 0.1875, -0.0625, 1.0, 100.0, -100.0, 15.9)] == [0, 0, 2, 0, 8, 127, -127, 127]


def test_int8_decoder_recovers_clean_codewords(oracle):
    n, k = 256, 128
    code = pb.Code(n, k, 0, 2.0)
    y = np.ones((3, n))                                             # all-zero codeword, no noise
    assert (oracle.decode_int8(n, 2, y, code.mask) == 0).all()
    rng = np.random.default_rng(0)
    yn = 1.0 + 0.3 * rng.standard_normal((50, n))
    assert (oracle.decode_int8(n, 2, yn, code.mask) == 0).all()
