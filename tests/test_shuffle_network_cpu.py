"""The ReverseShuffle of the packed encoder (polarcodes_b200/csrc/aux_kernels.cu: brev_stage / brev_stage16 / brev_stage8) is a
five-stage exchange network on 32 x 32 bit blocks: lane j holds word (j, c) of a frame, stage s swaps lane bit s with
position bit 4 - s, and lane t' then holds word (t', bitrev(c)) of the bit-reversed frame (Resources.cpp:148-158).
This restates the network stage by stage (with the kernel's masks, shifts and byte-permute selectors) and checks it
against the permutation itself for every code length the kernel takes.  No GPU needed."""
import numpy as np
import pytest

STAGES = ((0, 16, 0x0000FFFF), (1, 8, 0x00FF00FF), (2, 4, 0x0F0F0F0F), (3, 2, 0x33333333), (4, 1, 0x55555555))


def _brev(x, bits):
    return int(format(x, f"0{bits}b")[::-1], 2) if bits else 0


def _byte_perm(x, y, sel):                    # PRMT: selector nibble i picks byte 0-3 of x or 4-7 of y
    b = [(x >> (8 * i)) & 255 for i in range(4)] + [(y >> (8 * i)) & 255 for i in range(4)]
    return sum(b[(sel >> (4 * i)) & 7] << (8 * i) for i in range(4))


def _network(words, prmt):
    v = list(words)
    for s, d, m in STAGES:
        new = []
        for lane in range(32):
            x, y = v[lane], v[lane ^ (1 << s)]
            hi = (lane >> s) & 1
            if prmt and s == 0:
                new.append(_byte_perm(x, y, 0x3276 if hi else 0x5410))
            elif prmt and s == 1:
                new.append(_byte_perm(x, y, 0x3715 if hi else 0x6240))
            elif hi:
                new.append((x & ~m & 0xFFFFFFFF) | ((y & ~m & 0xFFFFFFFF) >> d))
            else:
                new.append((x & m) | (((y & m) << d) & 0xFFFFFFFF))
        v = new
    return v


@pytest.mark.parametrize("m", [10, 11, 12, 13, 15, 16])
@pytest.mark.parametrize("prmt", [False, True])
def test_exchange_network_is_the_bit_reversal_permutation(m, prmt):
    n, nw = 1 << m, (1 << m) // 32
    s_words, sb = nw // 32, m - 10
    rng = np.random.default_rng(m)
    bits = rng.integers(0, 2, n)
    pack = lambda b: [int(sum(int(b[32 * w + t]) << t for t in range(32))) for w in range(nw)]
    u = pack(bits)
    want = pack([bits[_brev(i, m)] for i in range(n)])
    got = [None] * nw
    for c in range(s_words):
        out = _network([u[lane * s_words + c] for lane in range(32)], prmt)
        for lane in range(32):
            got[lane * s_words + _brev(c, sb)] = out[lane]
    assert got == want
