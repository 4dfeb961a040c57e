"""Size-independent properties at the BASELINE.json sizes (the oracle is too slow there):
the tuned kernels against the reference-literal generic kernel on whole device-resident batches,
independence of the frames from their position in the batch (the tuned kernels decode several
frames per warp), and the noiseless round trip encode -> channel -> decode."""
import os

import numpy as np
import pytest

import polarcodes_b200 as pb

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

FULL = [  # name, n, k, crc, L, decoder, frames
    ("A  SC min-sum 1024", 1024, 512, 0, 1, pb.SCMINSUM, 1 << 17),
    ("B  SC exact 256", 256, 128, 0, 1, pb.SC, 1 << 16),
    ("C  SCL-8 1024 crc16", 1024, 512, 16, 8, pb.SCLMINSUM, 1 << 18),
    ("C  SCL-8 1024 crc0", 1024, 512, 0, 8, pb.SCLMINSUM, 1 << 17),
    ("   SCL-16 1024 crc8", 1024, 512, 8, 16, pb.SCLMINSUM, 1 << 15),
    ("conf.txt SCL-4 2048 crc16", 2048, 1024, 16, 4, pb.SCLMINSUM, 1 << 14),
    ("   SCL-2 1024 crc8", 1024, 512, 8, 2, pb.SCLMINSUM, (1 << 14) + 7),
    ("D  SCL-32 2048", 2048, 1024, 0, 32, pb.SCLMINSUM, 1 << 13),
    ("D  SCL-32 2048 crc16", 2048, 1024, 16, 32, pb.SCLMINSUM, (1 << 12) + 5),
    ("E  SC min-sum 32768", 32768, 16384, 0, 1, pb.SCMINSUM, 1 << 10),
]


def _handle(n, k, crc, L, dec, generic=False):
    if generic:
        os.environ["PCB_FORCE_GENERIC"] = "1"      # read when the handle is configured
    try:
        d = pb.Decoder(n, k, crc, L, dec, pb.STRICT, device=0)
    finally:
        os.environ.pop("PCB_FORCE_GENERIC", None)
    code = pb.Code(n, k, crc if dec == pb.SCLMINSUM else 0, 2.0)
    d.set_code(code.mask, code.sigma)
    return d


def _batch(d, n, B, snr=2.0, seed=77):
    y = torch.empty((B, n), dtype=torch.float32, device="cuda")
    tx = torch.empty((B, (n + 31) // 32), dtype=torch.int32, device="cuda")
    d.generate_dev(snr, B, y.data_ptr(), tx.data_ptr(), seed=seed, first_frame_id=0)
    return y, tx


def _decode(d, y):
    out = torch.zeros((y.shape[0], d.words), dtype=torch.int32, device="cuda")
    d.decode_dev(y.data_ptr(), True, out.data_ptr(), y.shape[0])
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("name,n,k,crc,L,dec,B", FULL, ids=[c[0].strip() for c in FULL])
def test_tuned_kernel_equals_generic_kernel_at_full_size(name, n, k, crc, L, dec, B):
    fast, slow = _handle(n, k, crc, L, dec), _handle(n, k, crc, L, dec, generic=True)
    assert fast.info()["blocks"] > 0 and slow.info()["blocks"] > 0
    y, _ = _batch(fast, n, B)
    a, b = _decode(fast, y), _decode(slow, y)
    if dec == pb.SC:   # tanh/atanh: same CUDA libm on both sides, still bit-exact between the two kernels
        assert (a != b).any(dim=1).sum().item() == 0
    else:
        bad = torch.nonzero((a != b).any(dim=1)).flatten()
        assert bad.numel() == 0, f"{name}: frames {bad[:8].tolist()} differ"


@pytest.mark.parametrize("n,k,crc,L,dec,B", [(1024, 512, 16, 8, pb.SCLMINSUM, 1 << 16), (2048, 1024, 0, 32, pb.SCLMINSUM, 1 << 11),
                                             (1024, 512, 0, 16, pb.SCLMINSUM, 1 << 13), (1024, 512, 0, 1, pb.SCMINSUM, 1 << 15),
                                             (2048, 1024, 16, 4, pb.SCLMINSUM, 1 << 12), (1024, 512, 0, 2, pb.SCLMINSUM, 1 << 12)])
def test_frames_do_not_depend_on_their_position(n, k, crc, L, dec, B):
    d = _handle(n, k, crc, L, dec)
    y, _ = _batch(d, n, B)
    ref = _decode(d, y)
    perm = torch.randperm(B, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
    assert torch.equal(_decode(d, y[perm].contiguous()), ref[perm])
    # a ragged prefix: the last warp holds fewer frames than it has room for
    cut = B - 3
    assert torch.equal(_decode(d, y[:cut].contiguous()), ref[:cut])


@pytest.mark.parametrize("n,k,crc,L,dec", [(1024, 512, 16, 8, pb.SCLMINSUM), (2048, 1024, 16, 32, pb.SCLMINSUM), (1024, 512, 16, 16, pb.SCLMINSUM),
                                           (1024, 512, 0, 1, pb.SCMINSUM), (256, 128, 0, 1, pb.SC), (32768, 16384, 0, 1, pb.SCMINSUM)])
def test_noiseless_round_trip(n, k, crc, L, dec):
    """info bits -> CRC -> frozen bits -> encode -> BPSK (40 dB: no sign flips) -> decode gives the info bits back.
    The list decoder returns the FIRST slot whose CRC passes (SCLDecoderMinsum.cpp:432-473, no metric sort), so a
    wrong path in a lower slot wins with probability 2^-16 per path even without noise (with CRC-8 that floor is
    15 % at L = 16: faithful behaviour): at most a couple of frames in 4096."""
    d = _handle(n, k, crc, L, dec)
    r = d.montecarlo(40.0, 1 << 12, seed=3)
    assert r["frames"] == 1 << 12
    if dec == pb.SCLMINSUM:
        assert r["frame_errors"] <= 2
    else:
        assert r["frame_errors"] == 0 and r["bit_errors"] == 0


def test_double_device_batch_equals_float_batch():
    """A device-resident double batch goes through the float pre-pass (the reference narrows first,
    SCLDecoderMinsum.cpp:406-408): same result as the float batch, including the frames the pre-pass
    sends to the generic kernel."""
    n, k, crc, L, B = 1024, 512, 16, 8, (1 << 18) + 6
    d = _handle(n, k, crc, L, pb.SCLMINSUM)
    y, _ = _batch(d, n, B)
    y[5, 17] = 1e-42                      # denormal: outside the tuned kernel's range
    y[B - 9, 3] = float("inf")
    ref = _decode(d, y)
    y64 = y.double()
    out = torch.zeros((B, d.words), dtype=torch.int32, device="cuda")
    for _ in range(2):
        d.decode_dev(y64.data_ptr(), False, out.data_ptr(), B)
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


def test_dev_entry_points_wait_for_the_callers_default_stream():
    """The handle's stream is non-blocking: a *_dev call is ordered after what the caller queued on the default stream
    (include/polar_b200.h, pcb_decode_batch_dev), so a batch still being written there is decoded complete."""
    n, k = 1024, 512
    d = _handle(n, k, 0, 1, pb.SCMINSUM)
    y, _ = _batch(d, n, 1 << 15)
    ref = _decode(d, y)
    for rep in range(3):
        late = torch.zeros_like(y)
        torch.cuda.synchronize()
        filler = torch.randn(1 << 26, device="cuda")
        for _ in range(4):
            filler = filler * 1.0001 + 1.0                      # a few ms of queued work in front of the copy
        late.copy_(y)
        out = torch.zeros((y.shape[0], d.words), dtype=torch.int32, device="cuda")
        d.decode_dev(late.data_ptr(), True, out.data_ptr(), y.shape[0])     # no synchronize in between
        assert torch.equal(out, ref)
