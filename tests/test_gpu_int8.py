"""PCB_PREC_INT8 (decoder 2, four 8-bit LLRs per register, polarcodes_b200/csrc/decode_sc8.cu) against its definition, the
CPU restatement oracle.decode_int8 (which is checked against the reference's Decodef_char / Decodeg_char in
tests/test_oracle_int8.py): bit-exact, for every size the mode takes, several scales, both input types, ragged batches,
saturating inputs and a mask with frozen ones.  The mode's FER is compared with the double decoder's statistically."""
import numpy as np
import pytest

import polarcodes_b200 as pb

pytestmark = pytest.mark.gpu


def _dec(n, k, q, mask=None, snr=2.0):
    d = pb.Decoder(n, k, 0, 1, pb.SCMINSUM, pb.int8_q(q), device=0)
    code = pb.Code(n, k, 0, snr)
    d.set_code(code.mask if mask is None else mask, code.sigma)
    return d, code


@pytest.mark.parametrize("n,k", [(64, 32), (256, 128), (1024, 512), (2048, 1024), (4096, 1000)])
@pytest.mark.parametrize("q", [1, 2, 4])
def test_int8_mode_equals_its_definition(oracle, n, k, q):
    d, code = _dec(n, k, q)
    rng = np.random.default_rng(n + q)
    for B in (1, 33, 300):
        y = 1.0 + code.sigma * rng.standard_normal((B, n))
        y[0, :8] = [0.0, -0.0, 1e9, -1e9, 0.0624, -0.0626, 127.5 / (2 << q), -1e-300]   # zeros, saturation, rounding ties
        y[0, 8:12] = [np.nan, -127.5 / (2 << q), np.inf, -np.inf]                      # the definition sends NaN to -127
        want = oracle.decode_int8(n, q, y, code.mask)
        got = d.decode(y)
        bad = np.flatnonzero((got != want).any(axis=1))
        assert bad.size == 0, f"n={n} q={q} B={B}: frames {bad[:8]} differ"
        yf = y.astype(np.float32)
        assert (d.decode(yf) == oracle.decode_int8(n, q, yf.astype(np.float64), code.mask)).all()


def test_int8_mode_with_frozen_ones(oracle):
    n, k, q = 512, 200, 2
    code = pb.Code(n, k, 0, 2.0)
    mask = np.array(code.mask).copy()
    mask[np.flatnonzero(mask == 0)[::5]] = 1
    d, _ = _dec(n, k, q, mask=mask)
    y = 1.0 + code.sigma * np.random.default_rng(1).standard_normal((257, n))
    assert (d.decode(y) == oracle.decode_int8(n, q, y, mask)).all()


def test_int8_mode_is_rejected_where_it_does_not_apply():
    with pytest.raises(pb.PolarError):
        pb.Decoder(1024, 512, 0, 8, pb.SCLMINSUM, pb.int8_q(2), device=0)       # decoder 2 only
    with pytest.raises(pb.PolarError):
        pb.Decoder(32768, 16384, 0, 1, pb.SCMINSUM, pb.int8_q(2), device=0)     # n <= 4096


def test_int8_mode_fer_stays_near_the_double_decoder():
    """(1024, 512) at 2.5 dB, 2^18 frames of the same Philox stream: the 8-bit decoder with the shipped scale (qFraction 2,
    conf.txt:9) loses a little; a coarse scale loses more; neither may be wildly off."""
    n, k, snr, frames = 1024, 512, 2.5, 1 << 18
    code = pb.Code(n, k, 0, snr)
    ref = pb.Decoder(n, k, 0, 1, pb.SCMINSUM, pb.STRICT, device=0)
    ref.set_code(code.mask, code.sigma)
    r0 = ref.montecarlo(snr, frames, seed=11)
    fer0 = r0["frame_errors"] / frames
    fers = {}
    for q in (1, 2, 3):
        d, _ = _dec(n, k, q, snr=snr)
        r = d.montecarlo(snr, frames, seed=11)
        fers[q] = r["frame_errors"] / frames
    assert fer0 > 1e-3
    assert fers[2] < 2.0 * fer0 and fers[3] < 2.0 * fer0 and fers[1] < 6.0 * fer0, (fer0, fers)
    assert min(fers.values()) > 0.5 * fer0, (fer0, fers)


def test_driver_runs_the_8_bit_mode_with_conf_txt_qfraction(tmp_path):
    """PCB_PREC=int8: conf.txt line 9 (qFraction, read and never used by the reference, PolarCodes.cpp:275) sets the scale;
    the counters are those of pcb_montecarlo on an INT8 handle with the same seed."""
    import os
    import subprocess
    host = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "polarcodes_b200", "host")
    subprocess.check_call(["make", "-s", "-C", host])
    (tmp_path / "conf.txt").write_text("1\n1024\n512\n50000\n2\n2\n1\n0\n3\n1\n2.5\n")
    env = dict(os.environ, PCB_PREC="int8", PCB_CSV=str(tmp_path / "o.csv"), PCB_SEED="77")
    r = subprocess.run([os.path.join(host, "polarcodes_driver")], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    row = open(tmp_path / "o.csv").read().strip().split("\n")[1].split(",")
    d, _ = _dec(1024, 512, 3, snr=2.5)
    want = d.montecarlo(2.5, 50000, seed=77)
    assert (int(row[1]), int(row[2])) == (want["frames"], want["frame_errors"])


@pytest.mark.parametrize("n,k,vs", [(1024, 512, 2), (1024, 512, 3), (1024, 512, 7), (2048, 1024, 4), (256, 128, 4), (64, 32, 3)])
def test_int8_mode_with_every_first_shared_memory_level(oracle, monkeypatch, n, k, vs):
    """PCB_8_VS moves levels between shared memory and the block's global slice (three inlined copies of the level op, one
    per address-space pair): the bits stay those of the definition."""
    monkeypatch.setenv("PCB_8_VS", str(vs))
    d, code = _dec(n, k, 2)
    rng = np.random.default_rng(n + vs)
    y = 1.0 + code.sigma * rng.standard_normal((500, n))
    assert (d.decode(y) == oracle.decode_int8(n, 2, y, code.mask)).all()


def test_int8_quantiser_as_a_pass_of_its_own(oracle, monkeypatch):
    """PCB_8_PREPASS=1 runs sc8_quantise_kernel in front of the decoder instead of the tile prologue: same bits."""
    monkeypatch.setenv("PCB_8_PREPASS", "1")
    for n, k in ((128, 64), (1024, 512)):
        d, code = _dec(n, k, 2)
        rng = np.random.default_rng(n)
        y = 1.0 + code.sigma * rng.standard_normal((333, n))
        y[3, :4] = [np.nan, 1e9, -1e9, 0.0]
        assert (d.decode(y) == oracle.decode_int8(n, 2, y, code.mask)).all()
        yf = y.astype(np.float32)
        assert (d.decode(yf) == oracle.decode_int8(n, 2, yf.astype(np.float64), code.mask)).all()
