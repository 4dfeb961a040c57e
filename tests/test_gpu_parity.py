"""GPU parity tests proper: the CUDA path (through the C ABI) against the committed golden
vectors (outputs of the unmodified reference) and against the CPU oracle on fresh seeded
inputs.  Bar: bit-exact for the encoder and the min-sum SC / SCL decoders."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

DECODE = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "sc*.npz")))


def _decoder(dec, n, k, crc, L, mask, sigma):
    import polarcodes_b200 as pb
    d = pb.Decoder(n, k, crc, L, dec)
    d.set_code(mask, sigma)
    return d


@pytest.mark.parametrize("name", DECODE)
def test_decode_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    dec, n, k, crc, L = (int(g[x]) for x in ("decoder", "n", "k", "crc", "L"))
    d = _decoder(dec, n, k, crc, L, g["mask"], float(g["sigma"]))
    sol = d.decode(g["y"])
    assert sol.shape == g["sol"].shape
    bad = np.nonzero((sol != g["sol"]).any(axis=1))[0]
    if dec == 1:
        # exact SC: CUDA tanh/atanh differ from glibc in ulps -> statistical parity only;
        # on these 41 frames a differing decision would need an LLR within ulps of zero.
        assert bad.size <= 1, f"{name}: frames {bad.tolist()} differ"
    else:
        assert bad.size == 0, f"{name}: frames {bad.tolist()} differ from the reference"


@pytest.mark.parametrize("name", [x for x in DECODE if x.startswith("scl")])
def test_decode_golden_f32_and_packed(name):
    import polarcodes_b200 as pb
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    dec, n, k, crc, L = (int(g[x]) for x in ("decoder", "n", "k", "crc", "L"))
    d = _decoder(dec, n, k, crc, L, g["mask"], float(g["sigma"]))
    yf = g["y"].astype(np.float32)   # SCLDecoderMinsum.cpp:406-408 makes this conversion itself
    assert (d.decode(yf) == g["sol"]).all()
    assert (pb.unpack_bits(d.decode_packed(yf), k + crc) == g["sol"]).all()


def test_encode_golden():
    import polarcodes_b200 as pb
    g = np.load(os.path.join(GOLDEN, "encode.npz"))
    for n in (2, 4, 8, 32, 64, 256, 1024, 2048):
        d = pb.Decoder(n, 1, 0, 1, pb.SCMINSUM)
        u = g[f"u{n}"]
        assert (d.encode(u, 0) == g[f"x{n}_nat"]).all(), n
        assert (d.encode(u, 1) == g[f"x{n}_shuf"]).all(), n


CASES = [  # decoder, n, k, crc, L, snr, frames
    (2, 1024, 512, 0, 1, 2.0, 64),
    (2, 64, 32, 0, 1, 1.0, 256),
    (2, 2, 1, 0, 1, 1.0, 64),
    (2, 8, 4, 0, 1, 1.0, 64),
    (2, 4096, 2048, 0, 1, 2.0, 8),
    (3, 1024, 512, 0, 8, 2.0, 64),
    (3, 1024, 512, 16, 8, 1.0, 64),
    (3, 1024, 512, 8, 8, 1.5, 64),
    (3, 2048, 1024, 16, 32, 2.0, 6),
    (3, 512, 256, 16, 16, 1.5, 32),
    (3, 256, 128, 8, 4, 1.5, 128),
    (3, 64, 32, 0, 2, 1.0, 256),
    (3, 32, 16, 0, 16, 1.0, 128),
    (3, 16, 8, 0, 4, 1.0, 128),
    (3, 8, 8, 0, 8, 1.0, 64),
    (3, 128, 64, 8, 1, 2.0, 128),
    (3, 4096, 2048, 16, 8, 2.0, 4),
    (3, 256, 128, 8, 8, 1.0, 256),
    (3, 512, 256, 16, 8, 1.5, 128),
    (3, 2048, 1024, 16, 8, 1.5, 32),
    (3, 1024, 256, 0, 8, 0.0, 64),
    (3, 1024, 768, 8, 8, 3.0, 64),
    # list 32 on the lane-per-path kernel (one frame per warp)
    (3, 2048, 1024, 0, 32, 2.0, 40),
    (3, 2048, 1024, 16, 32, 1.0, 40),
    (3, 1024, 512, 16, 32, 1.0, 64),
    (3, 512, 256, 8, 32, 1.5, 64),
    (3, 256, 128, 0, 32, 1.0, 96),
    (3, 1024, 128, 8, 32, 0.0, 33),
    # list 4 and 2 on the lane-per-path kernel (8 / 16 frames per warp)
    (3, 2048, 1024, 16, 4, 2.0, 67),
    (3, 1024, 512, 8, 4, 1.0, 131),
    (3, 512, 256, 0, 4, 1.5, 64),
    (3, 256, 128, 16, 4, 1.0, 250),
    (3, 1024, 512, 16, 2, 1.5, 97),
    (3, 2048, 1024, 0, 2, 2.0, 40),
    (3, 256, 128, 8, 2, 1.0, 129),
    (3, 512, 64, 8, 2, 0.0, 33),
    # list 16: two frames per warp
    (3, 1024, 512, 16, 16, 1.5, 65),
    (3, 2048, 1024, 0, 16, 2.0, 31),
    (3, 256, 128, 8, 16, 1.0, 128),
]


@pytest.mark.parametrize("dec,n,k,crc,L,snr,frames", CASES)
def test_cuda_equals_oracle(oracle, dec, n, k, crc, L, snr, frames):
    rng = np.random.default_rng(1000 + n + 7 * L + crc)
    sigma, inf, mask = oracle.make_code(n, k, crc, snr)
    # the all-zero codeword is a codeword: BPSK +1 plus noise exercises the same decode path
    xinf = rng.integers(0, 2, (frames, k + crc)).astype(np.int32)
    y = np.empty((frames, n))
    for f in range(frames):
        xi = oracle.append_crc(xinf[f], k, crc) if dec == 3 else xinf[f]
        x = oracle.add_frozen_bits(xi, n, inf)
        cw = oracle.encode(x, dec != 3)
        y[f] = (1 - 2 * cw) + sigma * rng.standard_normal(n)
    d = _decoder(dec, n, k, crc, L, mask, sigma)
    got = d.decode(y)
    want = oracle.decode(dec, n, L, k, crc, sigma, y, mask)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"frames {bad.tolist()} differ"


def test_quantised_inputs_force_ties(oracle):
    """Integer-grid LLRs make exact metric ties at the median common (pass 2,
    SCLDecoderMinsum.cpp:362-371) and exercise the slot-order discipline."""
    rng = np.random.default_rng(5)
    for n, k, crc, L in ((128, 64, 0, 8), (256, 128, 8, 8), (2048, 1024, 0, 8), (256, 128, 8, 4), (64, 32, 0, 16), (1024, 512, 16, 8), (256, 128, 0, 32),
                         (2048, 1024, 16, 32), (1024, 512, 0, 32), (512, 256, 0, 16), (1024, 512, 16, 16), (64, 32, 0, 1), (64, 32, 0, 2),
                         (1024, 512, 0, 4), (2048, 1024, 16, 4), (256, 128, 8, 2), (1024, 512, 16, 2), (512, 256, 0, 4)):
        sigma, inf, mask = oracle.make_code(n, k, crc, 2.0)
        y = rng.integers(-3, 4, (48, n)).astype(np.float64)
        d = _decoder(3, n, k, crc, L, mask, sigma)
        assert (d.decode(y) == oracle.decode(3, n, L, k, crc, sigma, y, mask)).all(), (n, L)


def test_frozen_value_one(oracle):
    """Encoder.cpp:43-55 lets frozen positions carry a value; the decoders must honour a 1."""
    rng = np.random.default_rng(9)
    n, k, L = 128, 64, 4
    sigma, inf, mask = oracle.make_code(n, k, 0, 2.0)
    mask = mask.copy()
    fz = np.nonzero(mask == 0)[0]
    mask[fz[::3]] = 1
    y = rng.standard_normal((32, n)) + 0.5
    for dec, LL in ((3, L), (2, 1)):
        d = _decoder(dec, n, k, 0, LL, mask, sigma)
        assert (d.decode(y) == oracle.decode(dec, n, LL, k, 0, sigma, y, mask)).all()


def test_large_batch_matches_small_batches(oracle):
    """Ragged / multi-chunk host batches: results are independent of batching."""
    rng = np.random.default_rng(3)
    n, k, crc, L = 256, 128, 8, 8
    sigma, inf, mask = oracle.make_code(n, k, crc, 1.5)
    y = (1 + sigma * rng.standard_normal((5000, n)))
    d = _decoder(3, n, k, crc, L, mask, sigma)
    full = d.decode(y)
    os.environ["PCB_CHUNK_FRAMES"] = "777"
    try:
        d2 = _decoder(3, n, k, crc, L, mask, sigma)
        parts = d2.decode(y)
    finally:
        del os.environ["PCB_CHUNK_FRAMES"]
    assert (full == parts).all()
    assert (full[:200] == oracle.decode(3, n, L, k, crc, sigma, y[:200], mask)).all()
    assert d.decode(y[:0]).shape == (0, k + crc)


def test_montecarlo_consistency(oracle):
    """On-device Philox frames: the device-side error counters equal what the oracle finds on
    the very same generated batch, and do not depend on how the frame range is split."""
    import polarcodes_b200 as pb
    n, k, crc, L, snr = 256, 128, 8, 4, 1.0
    code = pb.Code(n, k, crc, snr)
    d = _decoder(3, n, k, crc, L, code.mask, code.sigma)
    y, tx = d.generate(snr, 512, seed=77)
    for f in range(8):  # generated frames are self-consistent codewords
        assert (oracle.append_crc(tx[f], k, crc) == tx[f]).all()
    want = oracle.decode(3, n, L, k, crc, code.sigma, y.astype(np.float64), code.mask)
    errs = (want[:, :k] != tx[:, :k])
    r = d.montecarlo(snr, 512, seed=77)
    assert r["frames"] == 512
    assert r["frame_errors"] == int(errs.any(axis=1).sum())
    assert r["bit_errors"] == int(errs.sum())
    a = d.montecarlo(snr, 200, seed=77, first_frame_id=0)
    b = d.montecarlo(snr, 312, seed=77, first_frame_id=200)
    assert a["frame_errors"] + b["frame_errors"] == r["frame_errors"]
    assert a["bit_errors"] + b["bit_errors"] == r["bit_errors"]
    # (the noise statistics of the generator are tested in test_gpu_ref_parity.py::test_channel_noise_statistics)


@pytest.mark.parametrize("n,k,crc,L,snr,frames", [(1024, 512, 16, 8, 1.5, 40003), (1024, 512, 0, 8, 2.0, 9999), (2048, 1024, 16, 32, 1.5, 3001),
                                                  (2048, 1024, 16, 4, 2.0, 20000), (512, 256, 8, 16, 1.0, 7777), (256, 128, 8, 2, 1.0, 30001)])
def test_fused_montecarlo_equals_three_launch_path(n, k, crc, L, snr, frames):
    """PCB_MC_FUSED=1: pcb_montecarlo on the tuned list kernel generates the frames inside the warp that decodes
    them (no channel value in HBM) and scl_select_kernel counts; the default runs generate_kernel / list kernel /
    count_errors_kernel over a float batch in HBM.  Same Philox stream, same device functions: identical counters.
    PCB_TEST_FORCE_REDO_MOD pushes every 97th frame through the path for frames outside the tuned kernel's input range."""
    import polarcodes_b200 as pb
    code = pb.Code(n, k, crc, snr)
    d = _decoder(3, n, k, crc, L, code.mask, code.sigma)
    plain = d.montecarlo(snr, frames, seed=5, first_frame_id=123456789012)
    os.environ["PCB_MC_FUSED"] = "1"
    try:
        fused = d.montecarlo(snr, frames, seed=5, first_frame_id=123456789012)
        os.environ["PCB_TEST_FORCE_REDO_MOD"] = "97"
        forced = d.montecarlo(snr, frames, seed=5, first_frame_id=123456789012)
    finally:
        os.environ.pop("PCB_TEST_FORCE_REDO_MOD", None)
        del os.environ["PCB_MC_FUSED"]
    for key in ("frames", "frame_errors", "bit_errors"):
        assert fused[key] == plain[key] == forced[key], (key, fused, plain, forced)
    assert fused["frame_errors"] > 0 or snr >= 2.0


def test_sc_montecarlo_fer_matches_published():
    """pc.txt:3 (SC min-sum (1024,512), BEC construction, 2 dB): FER 0.1255 with 10 000 frames."""
    import polarcodes_b200 as pb
    code = pb.Code(1024, 512, 0, 2.0, zcomputer=1)
    d = _decoder(2, 1024, 512, 0, 1, code.mask, code.sigma)
    r = d.montecarlo(2.0, 20000, seed=11)
    fer = r["frame_errors"] / r["frames"]
    assert abs(fer - 0.1255) < 4 * np.sqrt(0.1255 * 0.8745 * (1 / 20000 + 1 / 10000)), fer


def test_out_of_range_frames_take_the_generic_kernel(oracle):
    """The tuned list-8 kernel hands frames with |y| outside {0} U [2^-40, 2^100] (or NaN/inf) to the
    generic kernel; a batch that mixes both kinds must still match the oracle frame by frame."""
    rng = np.random.default_rng(17)
    n, k, crc, L = 1024, 512, 16, 8
    sigma, inf, mask = oracle.make_code(n, k, crc, 2.0)
    y = 1 + sigma * rng.standard_normal((96, n))
    y[5] *= 1e-30            # products underflow in fp32
    y[11] *= 1e35            # sums overflow to inf, inf - inf = NaN
    y[12, 7] = 1e-20         # a single tiny value
    y[40] = 0.0
    y[41, ::3] = 0.0
    y[63, 100] = np.inf
    y[64, 200] = -3e38
    d = _decoder(3, n, k, crc, L, mask, sigma)
    got = d.decode(y)
    want = oracle.decode(3, n, L, k, crc, sigma, y, mask)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"frames {bad.tolist()} differ"
    assert (d.decode(y.astype(np.float32)) == oracle.decode(3, n, L, k, crc, sigma, y.astype(np.float32).astype(np.float64), mask)).all()


@pytest.mark.parametrize("frames", [1, 2, 3, 5, 7, 130])
def test_quad_kernel_ragged_batches(oracle, frames):
    """The list-8 kernel decodes four frames per warp: batches that are not a multiple of four
    (the last quad repeats the final frame and must not write it twice or out of bounds)."""
    rng = np.random.default_rng(100 + frames)
    n, k, crc, L = 1024, 512, 16, 8
    sigma, inf, mask = oracle.make_code(n, k, crc, 1.5)
    y = 1 + sigma * rng.standard_normal((frames, n))
    d = _decoder(3, n, k, crc, L, mask, sigma)
    want = oracle.decode(3, n, L, k, crc, sigma, y, mask)
    assert (d.decode(y) == want).all()
    import polarcodes_b200 as pb
    assert (pb.unpack_bits(d.decode_packed(y.astype(np.float32)), k + crc) ==
            oracle.decode(3, n, L, k, crc, sigma, y.astype(np.float32).astype(np.float64), mask)).all()


def test_quad_kernel_one_bad_frame_per_quad(oracle):
    """A quad with one out-of-range frame: the other three come from the quad kernel, the bad one
    from the generic kernel; also a quad whose four frames are all out of range."""
    rng = np.random.default_rng(23)
    n, k, crc, L = 1024, 512, 0, 8
    sigma, inf, mask = oracle.make_code(n, k, crc, 2.0)
    y = 1 + sigma * rng.standard_normal((24, n))
    y[1] *= 1e-30
    y[6, 3] = np.nan
    y[8:12] *= 1e36
    y[23, -1] = 1e-25
    d = _decoder(3, n, k, crc, L, mask, sigma)
    got, want = d.decode(y), oracle.decode(3, n, L, k, crc, sigma, y, mask)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"frames {bad.tolist()} differ"


@pytest.mark.parametrize("frac", ["0", "0.5", "1", "auto", "off"])
def test_host_pipeline_splits(oracle, frac):
    """pcb_decode_batch with double host buffers: every chunk is split into a head narrowed to float
    by the host thread pool and a tail narrowed by the device pre-pass.  Any split, ragged chunks and
    out-of-range frames on either side of the split give the reference's bits."""
    rng = np.random.default_rng(99)
    n, k, crc, L = 256, 128, 8, 8
    sigma, inf, mask = oracle.make_code(n, k, crc, 1.5)
    y = 1 + sigma * rng.standard_normal((203, n))
    y[3] *= 1e-30            # head of chunk 0
    y[36, 5] = np.inf        # tail of chunk 0
    y[100:104] *= 1e37
    y[202, 0] = 1e-44
    want = oracle.decode(3, n, L, k, crc, sigma, y, mask)
    env = {"PCB_CHUNK_FRAMES": "37"}
    if frac == "off":
        env["PCB_NO_HOST_NARROW"] = "1"
    elif frac != "auto":
        env["PCB_NARROW_FRAC"] = frac
    old = {key: os.environ.get(key) for key in env}
    os.environ.update(env)
    try:
        d = _decoder(3, n, k, crc, L, mask, sigma)
        got = d.decode(y)
        got2 = d.decode(y)       # second call: staging slots and rates re-used
    finally:
        for key, v in old.items():
            if v is None:
                del os.environ[key]
            else:
                os.environ[key] = v
    for g in (got, got2):
        bad = np.nonzero((g != want).any(axis=1))[0]
        assert bad.size == 0, f"frames {bad.tolist()} differ"
