"""The tile-per-block SC kernel (polarcodes_b200/csrc/decode_sct.cu: pruned tree walk, all-information subtrees decided from
the signs of their inputs, frames with a zero / vanishing LLR handed to the generic kernel) against the oracle and against
the reference-literal generic kernel, on the inputs that make the shortcut fail: quantised values (exact cancellations
inside the tree), zeros, denormals, huge values, ragged batches, frozen ones."""
import os

import numpy as np
import pytest

import polarcodes_b200 as pb

pytestmark = pytest.mark.gpu


def _dec(n, k, dec, prec=pb.STRICT, generic=False, mask=None, snr=2.0):
    if generic:
        os.environ["PCB_FORCE_GENERIC"] = "1"
    try:
        d = pb.Decoder(n, k, 0, 1, dec, prec, device=0)
    finally:
        os.environ.pop("PCB_FORCE_GENERIC", None)
    code = pb.Code(n, k, 0, snr)
    d.set_code(code.mask if mask is None else mask, code.sigma)
    return d, code


def _batch(n, B, sigma, seed, kind):
    rng = np.random.default_rng(seed)
    y = 1.0 + sigma * rng.standard_normal((B, n))
    if kind == "quantised":
        y = np.round(y * 4) / 4                              # multiples of 0.25: exact ties and zeros inside the tree
    elif kind == "coarse":
        y = np.round(y)                                      # -2 .. 4: zeros among the channel values themselves
    elif kind == "edge":
        idx = rng.integers(0, n, size=(B, 6))
        vals = np.array([0.0, -0.0, 1e-320, -1e-315, 1e300, -1e299])
        for b in range(B):
            y[b, idx[b]] = vals
    return y


@pytest.mark.parametrize("n,k", [(128, 64), (256, 128), (512, 300), (1024, 512), (2048, 1024)])
@pytest.mark.parametrize("kind", ["gauss", "quantised", "coarse", "edge"])
def test_tile_kernel_equals_generic_kernel_minsum(n, k, kind):
    fast, code = _dec(n, k, pb.SCMINSUM)
    slow, _ = _dec(n, k, pb.SCMINSUM, generic=True)
    assert fast.info()["warps_per_block"] in (1, 2, 4, 8) and fast.info()["first_smem_layer"] >= 2
    for B in (1, 33, 1000):
        y = _batch(n, B, code.sigma, n + B, kind)
        a, b = fast.decode(y), slow.decode(y)
        bad = np.flatnonzero((a != b).any(axis=1))
        assert bad.size == 0, f"{kind} n={n} B={B}: frames {bad[:8]} differ"


@pytest.mark.parametrize("kind", ["gauss", "quantised", "edge"])
def test_tile_kernel_equals_oracle(oracle, kind):
    n, k = 1024, 512
    fast, code = _dec(n, k, pb.SCMINSUM)
    y = _batch(n, 300, code.sigma, 7, kind)
    want = oracle.decode(2, n, 1, k, 0, code.sigma, y, code.mask)
    assert (fast.decode(y) == want).all()


def test_tile_kernel_fp32_mode_equals_generic_fp32():
    for n, k in ((256, 128), (1024, 512), (2048, 1024)):
        fast, code = _dec(n, k, pb.SCMINSUM, pb.FP32)
        slow, _ = _dec(n, k, pb.SCMINSUM, pb.FP32, generic=True)
        for kind in ("gauss", "quantised"):
            y = _batch(n, 777, code.sigma, n, kind)
            assert (fast.decode(y) == slow.decode(y)).all(), (n, kind)
            yf = y.astype(np.float32)
            assert (fast.decode(yf) == slow.decode(yf)).all(), (n, kind, "float input")


def test_tile_kernel_with_frozen_ones_and_odd_rates(oracle):
    n = 512
    for k in (8, 100, 400, 504):
        code = pb.Code(n, k, 0, 2.0)
        mask = np.array(code.mask).copy()
        frozen = np.flatnonzero(mask == 0)
        mask[frozen[::3]] = 1
        fast, _ = _dec(n, k, pb.SCMINSUM, mask=mask)
        y = _batch(n, 200, code.sigma, k, "gauss")
        want = oracle.decode(2, n, 1, k, 0, code.sigma, y, mask)
        assert (fast.decode(y) == want).all(), k


def test_exact_sc_tile_kernel_against_generic_kernel():
    """tanh/atanh of CUDA's libm in both kernels: the pruned walk must give the same bits as the full walk except where a
    subtree's shortcut is not the literal computation (never on these inputs)."""
    n, k = 256, 128
    fast, code = _dec(n, k, pb.SC)
    slow, _ = _dec(n, k, pb.SC, generic=True)
    y = _batch(n, 4096, code.sigma, 3, "gauss")
    a, b = fast.decode(y), slow.decode(y)
    assert ((a != b).any(axis=1)).sum() <= 2


@pytest.mark.parametrize("n,k,prec,W,VS", [
    (1024, 512, pb.STRICT, 1, 2), (1024, 512, pb.STRICT, 2, 3), (1024, 512, pb.STRICT, 4, 5), (1024, 512, pb.STRICT, 8, 4),
    (1024, 512, pb.FP32, 1, 5), (1024, 512, pb.FP32, 2, 2), (1024, 512, pb.FP32, 8, 6), (256, 128, pb.STRICT, 2, 2),
    (256, 128, pb.FP32, 4, 3), (2048, 1024, pb.STRICT, 2, 6), (4096, 2048, pb.STRICT, 1, 6), (4096, 2048, pb.STRICT, 2, 4),
    (4096, 2048, pb.FP32, 8, 6), (8192, 4096, pb.STRICT, 4, 7)])
def test_every_launch_shape_decodes_the_same(monkeypatch, n, k, prec, W, VS):
    """Warps per tile and the first shared-memory level select different instantiations of the kernel (levels addressed as
    shared memory, as global memory, or through generic pointers; heap in shared memory or in the block's global slice):
    each one must give the bits of the reference-literal kernel."""
    monkeypatch.setenv("PCB_T_W", str(W))
    monkeypatch.setenv("PCB_T_VS", str(VS))
    try:
        fast, code = _dec(n, k, pb.SCMINSUM, prec)
    except pb.PolarError:
        pytest.skip("shape does not fit an SM")
    monkeypatch.delenv("PCB_T_W")
    monkeypatch.delenv("PCB_T_VS")
    i = fast.info()
    if (i["warps_per_block"], i["first_smem_layer"]) != (W, VS):
        pytest.skip(f"shape not honoured for this code: {i}")
    slow, _ = _dec(n, k, pb.SCMINSUM, prec, generic=True)
    for kind, B in (("gauss", 700), ("quantised", 65)):
        y = _batch(n, B, code.sigma, n + W + VS, kind)
        bad = np.flatnonzero((fast.decode(y) != slow.decode(y)).any(axis=1))
        assert bad.size == 0, f"n={n} W={W} VS={VS} {kind}: frames {bad[:8]} differ"
