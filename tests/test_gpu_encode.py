"""Packed encoder entry points (pcb_encode_batch_packed / pcb_encode_batch_dev) against the int-per-bit pcb_encode_batch, which
the golden vectors of tests/test_gpu_parity.py pin to the reference (Encoder.cpp:57-83, Resources.cpp:148-158)."""
import numpy as np
import pytest

import polarcodes_b200 as pb

pytestmark = pytest.mark.gpu


def _pack(bits):
    B, n = bits.shape
    nw = max(n // 32, 1)
    out = np.zeros((B, nw), np.uint32)
    for t in range(n):
        out[:, t >> 5] |= (bits[:, t].astype(np.uint32) & 1) << np.uint32(t & 31)
    return out


@pytest.mark.parametrize("n", [8, 32, 64, 512, 1024, 2048, 4096, 16384, 32768, 65536])
@pytest.mark.parametrize("shuffle", [0, 1])
def test_packed_encoder_equals_int_per_bit_encoder(n, shuffle):
    d = pb.Decoder(n, max(n // 2, 1), 0, 1, pb.SCMINSUM, device=0)
    rng = np.random.default_rng(n + shuffle)
    for B in (1, 5, 333):
        u = rng.integers(0, 2, size=(B, n)).astype(np.int32)
        want = _pack(d.encode(u, shuffle))
        got = d.encode_packed(_pack(u), shuffle)
        assert (got == want).all(), (n, shuffle, B)


def test_device_encoder_equals_host_packed_encoder():
    torch = pytest.importorskip("torch")
    n, B = 2048, 10000
    d = pb.Decoder(n, 1024, 0, 1, pb.SCMINSUM, device=0)
    u = torch.randint(-2 ** 31, 2 ** 31 - 1, (B, n // 32), dtype=torch.int32, device="cuda")
    x = torch.zeros_like(u)
    for shuffle in (0, 1):
        d.encode_dev(u.data_ptr(), x.data_ptr(), shuffle, B)
        torch.cuda.synchronize()
        want = d.encode_packed(u.cpu().numpy().view(np.uint32), shuffle)
        assert (x.cpu().numpy().view(np.uint32) == want).all()
        # the transform is an involution in natural order
        if not shuffle:
            y = torch.zeros_like(u)
            d.encode_dev(x.data_ptr(), y.data_ptr(), 0, B)
            torch.cuda.synchronize()
            assert torch.equal(y, u)


@pytest.mark.parametrize("n", [1024, 2048, 4096, 8192, 16384, 32768])
def test_register_and_shared_memory_encoders_agree(n, monkeypatch):
    """n = 1024 .. 32768 run a register-resident kernel; PCB_ENC_SMEM=1 selects the shared-memory one all other sizes use."""
    d = pb.Decoder(n, n // 2, 0, 1, pb.SCMINSUM, device=0)
    rng = np.random.default_rng(n)
    u = rng.integers(0, 2 ** 32, size=(1001, n // 32), dtype=np.uint64).astype(np.uint32)
    for shuffle in (0, 1):
        fast = d.encode_packed(u, shuffle)
        monkeypatch.setenv("PCB_ENC_SMEM", "1")
        slow = d.encode_packed(u, shuffle)
        monkeypatch.delenv("PCB_ENC_SMEM")
        assert (fast == slow).all(), (n, shuffle)
