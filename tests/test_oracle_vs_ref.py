"""The C oracle against the compiled reference on fresh seeded inputs (needs oracle/_ref;
skipped where it is not built).  CPU only, sized to run in well under a minute."""
import numpy as np
import pytest

CASES = [  # decoder, n, k, crc, L, snr, frames
    (2, 1024, 512, 0, 1, 2.0, 60),
    (2, 64, 32, 0, 1, 1.0, 200),
    (2, 2, 1, 0, 1, 1.0, 50),
    (1, 256, 128, 0, 1, 2.0, 100),
    (3, 1024, 512, 0, 8, 2.0, 30),
    (3, 1024, 512, 16, 8, 1.0, 30),
    (3, 1024, 512, 8, 8, 1.5, 30),
    (3, 2048, 1024, 16, 32, 2.0, 4),
    (3, 256, 128, 8, 4, 1.5, 100),
    (3, 64, 32, 0, 2, 1.0, 300),
    (3, 32, 16, 0, 16, 1.0, 100),
    (3, 128, 64, 8, 1, 2.0, 100),
]


@pytest.mark.ref
@pytest.mark.parametrize("dec,n,k,crc,L,snr,frames", CASES)
def test_port_equals_reference(oracle, ref, dec, n, k, crc, L, snr, frames):
    ref.seed(99 + n + L)
    sigma, inf, mask = ref.make_code(n, k, crc, snr)
    so, io, mo = oracle.make_code(n, k, crc, snr)
    assert sigma == so and (inf == io).all() and (mask == mo).all()
    xinf, x, cw, y = ref.gen_frames(n, k, crc, inf, sigma, dec != 3, frames)
    a = ref.decode(dec, n, L, k, crc, sigma, y, mask)
    b = oracle.decode(dec, n, L, k, crc, sigma, y, mask)
    assert (a == b).all()


@pytest.mark.ref
def test_quantised_inputs_force_ties(oracle, ref):
    """Integer-grid LLRs make exact path-metric ties common: pass 2 of the median select
    (SCLDecoderMinsum.cpp:362-371) and the slot-order discipline are load-bearing here."""
    rng = np.random.default_rng(5)
    for n, k, crc, L in ((128, 64, 0, 8), (256, 128, 8, 4), (64, 32, 0, 16), (1024, 512, 16, 8)):
        sigma, inf, mask = ref.make_code(n, k, crc, 2.0)
        y = rng.integers(-3, 4, (40, n)).astype(np.float64)
        assert (ref.decode(3, n, L, k, crc, sigma, y, mask) == oracle.decode(3, n, L, k, crc, sigma, y, mask)).all()
