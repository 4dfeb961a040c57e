"""pcb_decode_batch_dev_noalloc (include/polar_b200.h): the decode entry point that allocates nothing —
its kernel's LLR scratch is a static device array — against pcb_decode_batch_dev on the same device
batches, bit for bit (integer output: the tolerance of every decode test here is zero), on two calls in
a row (the array keeps the previous call's contents), through the C ABI.  Collected last on purpose."""
import pytest

import polarcodes_b200 as pb

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

CASES = [  # name, n, k, crc, L, decoder, frames of call 1, frames of call 2, double input
    ("SCL-8 1024 crc16", 1024, 512, 16, 8, pb.SCLMINSUM, 4096, 1531, False),
    ("SCL-8 1024 crc0 double", 1024, 512, 0, 8, pb.SCLMINSUM, 2048, 2048, True),
    ("SCL-32 2048 crc16", 2048, 1024, 16, 32, pb.SCLMINSUM, 512, 300, False),
    ("SC min-sum 1024 fp64", 1024, 512, 0, 1, pb.SCMINSUM, 8192, 4099, True),
    ("SC min-sum 32768 fp64", 32768, 16384, 0, 1, pb.SCMINSUM, 96, 64, True),
    ("SC exact 256", 256, 128, 0, 1, pb.SC, 4096, 4096, False),
]


def _handle(n, k, crc, L, dec, snr=2.0):
    d = pb.Decoder(n, k, crc, L, dec, pb.STRICT, device=0)
    code = pb.Code(n, k, crc if dec == pb.SCLMINSUM else 0, snr)
    d.set_code(code.mask, code.sigma)
    return d


def _batch(d, n, B, seed, as_double):
    y = torch.empty((B, n), dtype=torch.float32, device="cuda")
    d.generate_dev(1.5, B, y.data_ptr(), None, seed=seed, first_frame_id=0)
    torch.cuda.synchronize()
    return y.double() if as_double else y


@pytest.mark.parametrize("name,n,k,crc,L,dec,B1,B2,dbl", CASES, ids=[c[0] for c in CASES])
def test_noalloc_equals_decode_batch_dev_twice(name, n, k, crc, L, dec, B1, B2, dbl):
    d = _handle(n, k, crc, L, dec)
    batches = [_batch(d, n, B1, 1234, dbl), _batch(d, n, B2, 99, dbl)]
    want = []
    for y in batches:
        out = torch.zeros((y.shape[0], d.words), dtype=torch.int32, device="cuda")
        d.decode_dev(y.data_ptr(), not dbl, out.data_ptr(), y.shape[0])
        want.append(out)
    got = [torch.full((y.shape[0], d.words), -1, dtype=torch.int32, device="cuda") for y in batches]
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for y, out in zip(batches, got):                       # two calls in a row on the shared static array
        d.decode_dev_noalloc(y.data_ptr(), not dbl, out.data_ptr(), y.shape[0])
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    assert free_after == free_before, "the no-allocation entry point changed the device's free memory"
    for i, (a, b) in enumerate(zip(got, want)):
        bad = (a != b).any(dim=1).sum().item()
        assert bad == 0, f"{name}: call {i + 1}: {bad} of {a.shape[0]} frames differ from pcb_decode_batch_dev"
    # and the first batch again, after the array has held another batch's layers
    again = torch.full_like(got[0], -1)
    d.decode_dev_noalloc(batches[0].data_ptr(), not dbl, again.data_ptr(), batches[0].shape[0])
    torch.cuda.synchronize()
    assert torch.equal(again, want[0])
    d.close()


def test_noalloc_two_handles_share_the_array_in_turn():
    """Different codes, one device, alternating calls: every call rewrites what it reads."""
    a, b = _handle(1024, 512, 16, 8, pb.SCLMINSUM), _handle(2048, 1024, 0, 4, pb.SCLMINSUM)
    ya, yb = _batch(a, 1024, 1024, 5, False), _batch(b, 2048, 777, 6, False)
    wa = torch.zeros((1024, a.words), dtype=torch.int32, device="cuda")
    wb = torch.zeros((777, b.words), dtype=torch.int32, device="cuda")
    a.decode_dev(ya.data_ptr(), True, wa.data_ptr(), 1024)
    b.decode_dev(yb.data_ptr(), True, wb.data_ptr(), 777)
    for _ in range(2):
        ga, gb = torch.zeros_like(wa), torch.zeros_like(wb)
        a.decode_dev_noalloc(ya.data_ptr(), True, ga.data_ptr(), 1024)
        b.decode_dev_noalloc(yb.data_ptr(), True, gb.data_ptr(), 777)
        torch.cuda.synchronize()
        assert torch.equal(ga, wa) and torch.equal(gb, wb)
    a.close()
    b.close()


def test_noalloc_rejects_sizes_beyond_the_maximum():
    # more frames than one call takes (checked before anything is launched or read)
    d = _handle(256, 128, 0, 1, pb.SCMINSUM)
    y = torch.zeros((1, 256), dtype=torch.float32, device="cuda")
    out = torch.zeros((1, d.words), dtype=torch.int32, device="cuda")
    with pytest.raises(pb.PolarError, match="exceeds the limit"):
        d.decode_dev_noalloc(y.data_ptr(), True, out.data_ptr(), (1 << 20) + 1)
    with pytest.raises(pb.PolarError, match="null"):
        d.decode_dev_noalloc(None, True, out.data_ptr(), 1)
    d.close()
    # the largest list code a handle can hold (n * L = 2^19: 2 MiB of LLR scratch per one-warp block) still fits the array
    d = pb.Decoder(32768, 16384, 0, 16, pb.SCLMINSUM, pb.STRICT, device=0)
    code = pb.Code(32768, 16384, 0, 2.0)
    d.set_code(code.mask, code.sigma)
    yb = _batch(d, 32768, 24, 3, False)
    want, got = (torch.zeros((24, d.words), dtype=torch.int32, device="cuda") for _ in range(2))
    d.decode_dev(yb.data_ptr(), True, want.data_ptr(), 24)
    d.decode_dev_noalloc(yb.data_ptr(), True, got.data_ptr(), 24)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    d.close()
