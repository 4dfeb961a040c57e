"""GPU tests of the C++ host side: the reference-signature shim (B = 1 per call) reproduces the
golden outputs of the reference, and the conf.txt driver writes a reference-format result file
whose FER agrees with the published pc.txt numbers."""
import glob
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu
HOST = os.path.join(ROOT, "polarcodes_b200", "host")

SHIM_PROG = r"""
#include <cstdio>
#include <vector>
#include "polar_b200_shim.h"
// argv: decoder n k crc L frames ; stdin: mask[n] ints, sigma, y[frames][n] doubles (binary)
int main(int argc, char** argv) {
    int dec = atoi(argv[1]), n = atoi(argv[2]), k = atoi(argv[3]), crc = atoi(argv[4]), L = atoi(argv[5]), F = atoi(argv[6]);
    std::vector<int> mask(n); double sigma;
    fread(mask.data(), 4, n, stdin); fread(&sigma, 8, 1, stdin);
    std::vector<double> y(n); std::vector<int> sol(n);
    if (dec == 3) SCLDecoderMinsum::InitializeDataStructures(n, L, k, crc);
    int stride = dec == 3 ? k + crc : n;
    for (int f = 0; f < F; f++) {
        fread(y.data(), 8, n, stdin);
        if (dec == 1) SCDecoder::Decode(n, y.data(), nullptr, nullptr, sol.data(), mask, sigma);
        else if (dec == 2) SCDecoderMinsum::Decode(n, y.data(), nullptr, nullptr, sol.data(), mask);
        else SCLDecoderMinsum::Decode(n, L, y.data(), nullptr, sol.data(), mask, 2);
        fwrite(sol.data(), 4, stride, stdout);
    }
    // Encode through the shim: result in encTree[0]
    std::vector<int> u(n, 0), out(n); u[n - 1] = 1; int* tree[1] = {out.data()};
    Encode(n, u.data(), tree, false);
    for (int i = 0; i < n; i++) if (out[i] != 1) return 3;   // last row of F^(x)m is all ones
    return 0;
}
"""


@pytest.fixture(scope="module")
def shim_exe(tmp_path_factory):
    d = tmp_path_factory.mktemp("shim")
    (d / "prog.cpp").write_text(SHIM_PROG)
    lib = os.path.join(ROOT, "polarcodes_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-w", "-I", HOST, str(d / "prog.cpp"), "-L", lib, "-lpolar_b200",
                           f"-Wl,-rpath,{lib}", "-Wl,-rpath,/usr/local/cuda/lib64", "-o", str(d / "prog")])
    return str(d / "prog")


@pytest.mark.parametrize("name", ["scl8_1024_512_crc16", "scl4_256_128_crc8", "scms_1024_512", "scl2_64_32_crc0"])
def test_shim_matches_golden(shim_exe, name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    dec, n, k, crc, L = (int(g[x]) for x in ("decoder", "n", "k", "crc", "L"))
    F = min(12, g["y"].shape[0])
    blob = g["mask"].astype(np.int32).tobytes() + np.float64(g["sigma"]).tobytes() + g["y"][:F].astype(np.float64).tobytes()
    r = subprocess.run([shim_exe] + [str(v) for v in (dec, n, k, crc, L, F)], input=blob, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()
    sol = np.frombuffer(r.stdout, np.int32).reshape(F, -1)
    assert (sol == g["sol"][:F]).all()


def test_driver_reproduces_pc_txt(tmp_path):
    """pc.txt:1-9 col 2 (SC min-sum (1024,512), BEC construction): 0.3611 @1.5 dB, 0.1255 @2 dB, 0.0351 @2.5 dB."""
    subprocess.check_call(["make", "-s", "-C", HOST])
    (tmp_path / "conf.txt").write_text("1\n1024\n512\n40000\n2\n1\n1\n0\n2\n3\n1.5 2 2.5\n")
    env = dict(os.environ, PCB_CSV=str(tmp_path / "out.csv"))
    r = subprocess.run([os.path.join(HOST, "polarcodes_driver")], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    files = glob.glob(str(tmp_path / "SC_1_1024_512_L1_CRC0_*.txt"))
    assert len(files) == 1
    lines = open(files[0]).read().split("\n")
    int(lines[0])   # elapsed ticks, PolarCodes.cpp:304
    got = {float(a): float(b) for a, b in (ln.split("\t") for ln in lines[1:] if ln)}
    want = {1.5: 0.3611, 2.0: 0.1255, 2.5: 0.0351}
    for snr, p in want.items():
        tol = 4 * np.sqrt(p * (1 - p) * (1 / 40000 + 1 / 10000))
        assert abs(got[snr] - p) < tol, (snr, got[snr], p)
    csv = open(tmp_path / "out.csv").read().strip().split("\n")
    assert csv[0].startswith("snr_db,frames,frame_errors,bit_errors") and len(csv) == 4


def test_driver_scl_crc16_and_throughput_mode(tmp_path):
    subprocess.check_call(["make", "-s", "-C", HOST])
    (tmp_path / "conf.txt").write_text("2\n1024\n512\n20000\n3\n2\n8\n16\n2\n1\n2\n")
    r = subprocess.run([os.path.join(HOST, "polarcodes_driver")], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    files = glob.glob(str(tmp_path / "SC_2_1024_512_L8_CRC16_*.txt"))
    assert len(files) == 1
    ticks = float(open(files[0]).read().split("\n")[1].split("\t")[1])
    assert ticks > 0   # decode time of 20000 frames in clock ticks


def _csv_rows(path):
    rows = open(path).read().strip().split("\n")[1:]
    return [tuple(int(v) for v in r.split(",")[1:4]) for r in rows]   # frames, frame errors, bit errors


def test_driver_checkpoint_resume_and_error_budget(tmp_path):
    """an interrupted run resumed from its checkpoint ends with exactly the counters
    of an uninterrupted one (Philox is keyed by the global frame id); PCB_MAX_ERRORS stops a point early."""
    subprocess.check_call(["make", "-s", "-C", HOST])
    exe = os.path.join(HOST, "polarcodes_driver")
    (tmp_path / "conf.txt").write_text("1\n1024\n512\n30000\n2\n2\n1\n0\n2\n2\n1.5 2.5\n")
    base = dict(os.environ, PCB_SLICE="8192")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, env=dict(base, PCB_CSV=str(tmp_path / "a.csv")))
    assert r.returncode == 0, r.stderr
    want = _csv_rows(tmp_path / "a.csv")
    assert [w[0] for w in want] == [30000, 30000]
    ck = dict(base, PCB_CKPT=str(tmp_path / "run.ckpt"), PCB_CSV=str(tmp_path / "b.csv"))
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, env=dict(ck, PCB_STOP_AFTER_SLICES="2"))
    assert r.returncode == 3 and os.path.exists(tmp_path / "run.ckpt")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, env=dict(ck, PCB_STOP_AFTER_SLICES="3"))
    assert r.returncode == 3                      # second interruption, inside the second SNR point
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, env=ck)
    assert r.returncode == 0 and "Resuming" in r.stdout, r.stderr
    assert _csv_rows(tmp_path / "b.csv") == want
    # a checkpoint of another configuration is ignored
    (tmp_path / "conf.txt").write_text("1\n1024\n512\n30000\n2\n2\n1\n0\n2\n1\n1.5\n")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, env=dict(ck, PCB_CSV=str(tmp_path / "c.csv")))
    assert r.returncode == 0 and "another configuration" in r.stderr
    assert _csv_rows(tmp_path / "c.csv") == want[:1]
    # error budget
    os.remove(tmp_path / "run.ckpt")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True,
                       env=dict(base, PCB_MAX_ERRORS="100", PCB_CSV=str(tmp_path / "d.csv")))
    assert r.returncode == 0, r.stderr
    (frames, ferr, _), = _csv_rows(tmp_path / "d.csv")
    assert frames == 8192 and ferr >= 100       # FER ~ 0.36 at 1.5 dB: the first slice is enough


def _result_file(tmp_path, pattern):
    files = glob.glob(str(tmp_path / pattern))
    assert len(files) == 1, files
    lines = open(files[0]).read().split("\n")
    return {float(a): float(b) for a, b in (ln.split("\t") for ln in lines[1:] if ln)}


def test_reference_main_runs_on_the_gpu_through_the_shim(tmp_path):
    """oracle/_ref/PolarCodes_shim is the reference's own main() with the one edit of INTEGRATION.md section 1,
    built where /root/reference exists: its unchanged Monte-Carlo loop (PolarCodes.cpp:94-248, mt19937 frames)
    calls Encode / SCDecoderMinsum::Decode / SCLDecoderMinsum::Decode, which run on the GPU one frame per call.
    Same conf.txt through the batched driver: the two FERs agree within binomial bounds."""
    exe = os.path.join(ROOT, "oracle", "_ref", "PolarCodes_shim")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/PolarCodes_shim not built (reference sources absent)")
    subprocess.check_call(["make", "-s", "-C", HOST])
    for conf, pat, frames in (("1\n1024\n512\n2000\n2\n1\n1\n0\n2\n1\n1.5\n", "SC_1_1024_512_L1_CRC0_*.txt", 2000),
                              ("1\n256\n128\n2000\n3\n2\n4\n8\n2\n1\n1.0\n", "SC_1_256_128_L4_CRC8_*.txt", 2000)):
        a, b = tmp_path / ("shim" + pat[5:9]), tmp_path / ("drv" + pat[5:9])
        a.mkdir(); b.mkdir()
        (a / "conf.txt").write_text(conf)
        (b / "conf.txt").write_text(conf.replace("\n2000\n", "\n40000\n"))
        r = subprocess.run([exe], cwd=a, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        assert "Running..." in r.stdout
        r2 = subprocess.run([os.path.join(HOST, "polarcodes_driver")], cwd=b, capture_output=True, text=True)
        assert r2.returncode == 0, r2.stderr
        (snr, fer_shim), = _result_file(a, pat).items()
        (snr2, fer_drv), = _result_file(b, pat).items()
        assert snr == snr2
        p = (fer_shim * frames + fer_drv * 40000) / (frames + 40000)
        assert abs(fer_shim - fer_drv) < 4 * np.sqrt(max(p * (1 - p), 1e-6) * (1 / frames + 1 / 40000)), (fer_shim, fer_drv)


def test_driver_reproduces_both_pc_txt_columns(tmp_path):
    """All nine points of pc.txt:1-9, column 2 (SC min-sum) and column 3 (exact SC, pc.plt:10), (1024,512),
    BEC construction; the published table was made with 10 000 frames per point."""
    subprocess.check_call(["make", "-s", "-C", HOST])
    snrs = [1, 1.5, 2, 2.5, 3, 3.5, 4, 4.5, 5]
    col2 = [0.7283, 0.3611, 0.1255, 0.0351, 0.0084, 0.0018, 0.0004, 0.0001, 0.0]
    col3 = [0.7054, 0.3662, 0.1251, 0.0348, 0.0099, 0.0018, 0.0007, 0.0001, 0.0]
    frames = 200000
    for dec, want in ((2, col2), (1, col3)):
        d = tmp_path / f"dec{dec}"
        d.mkdir()
        (d / "conf.txt").write_text(f"1\n1024\n512\n{frames}\n{dec}\n1\n1\n0\n2\n9\n" + " ".join(str(s) for s in snrs) + "\n")
        r = subprocess.run([os.path.join(HOST, "polarcodes_driver")], cwd=d, capture_output=True, text=True,
                           env=dict(os.environ, PCB_CSV=str(d / "out.csv")))
        assert r.returncode == 0, r.stderr
        got = _result_file(d, "SC_1_1024_512_L1_CRC0_*.txt")
        for snr, p in zip(snrs, want):
            # pc.txt is rounded to 1e-4 and has 10 000 frames: binomial bounds of both samples plus the rounding
            pp = max(p, got[snr], 1e-4)
            tol = 4 * np.sqrt(pp * (1 - pp) * (1 / frames + 1 / 10000)) + 1e-4
            assert abs(got[snr] - p) < tol, (dec, snr, got[snr], p)
        head = open(d / "out.csv").readline().strip().split(",")
        assert head[-2:] == ["lane_ops_per_frame", "alu_roofline_frac"]
        row = open(d / "out.csv").read().strip().split("\n")[1].split(",")
        assert float(row[-2]) == 30880.0 and 0 < float(row[-1]) < 1       # config A's algorithmic work


def test_driver_sc_decoder_with_crc_in_conf(tmp_path):
    """conf.txt with CRC = 16 and decoder 2: the code has k + 16 information positions (PolarCodes.cpp:125,136-155);
    FER against the reference's own loop."""
    from oracle.pyoracle import Ref
    if not Ref.available():
        pytest.skip("oracle/_ref not built")
    subprocess.check_call(["make", "-s", "-C", HOST])
    (tmp_path / "conf.txt").write_text("1\n1024\n512\n100000\n2\n2\n1\n16\n2\n1\n2.0\n")
    r = subprocess.run([os.path.join(HOST, "polarcodes_driver")], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    (snr, fer), = _result_file(tmp_path, "SC_1_1024_512_L1_CRC16_*.txt").items()
    ref = Ref()
    ref.seed(7)
    p_ref = ref.fer_point(1024, 512, 16, 1, 2, 2, 2.0, 20000) / 20000
    assert abs(fer - p_ref) < 4 * np.sqrt(max(p_ref * (1 - p_ref), 1e-6) * (1 / 100000 + 1 / 20000)), (fer, p_ref)


def _spawn_ranks(exe, cwd, world, env_extra, port):
    procs = []
    for r in range(world):
        env = dict(os.environ, WORLD_SIZE=str(world), RANK=str(r), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), PCB_NCCL_ID_FILE=str(cwd / f"nccl_id_{port}"), **env_extra)
        procs.append(subprocess.Popen([exe], cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=600) for p in procs]
    return procs, outs


def test_driver_multi_process_nccl_equals_one_gpu(tmp_path):
    """One process per GPU, ncclAllReduce(SUM, uint64[4]) of the counters after every slice (C++, in the product):
    the counters equal those of the one-process run, and the cross-rank error budget stops every rank at the
    same slice.  Needs two visible GPUs (skipped on a one-GPU box; bench.py covers the NCCL path at N > 1)."""
    import polarcodes_b200 as pb
    if pb.lib().pcb_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    subprocess.check_call(["make", "-s", "-C", HOST])
    exe = os.path.join(HOST, "polarcodes_driver")
    (tmp_path / "conf.txt").write_text("1\n1024\n512\n50000\n3\n2\n8\n16\n2\n2\n1.0 2.0\n")
    base = dict(PCB_SLICE="16384")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, env=dict(os.environ, PCB_CSV=str(tmp_path / "one.csv"), **base))
    assert r.returncode == 0, r.stderr
    procs, outs = _spawn_ranks(exe, tmp_path, 2, dict(PCB_CSV=str(tmp_path / "two.csv"), **base), 29611)
    assert all(p.returncode == 0 for p in procs), outs
    assert outs[1][0] == ""                                   # rank 1 is silent
    assert _csv_rows(tmp_path / "two.csv") == _csv_rows(tmp_path / "one.csv")
    procs, outs = _spawn_ranks(exe, tmp_path, 2, dict(PCB_CSV=str(tmp_path / "budget.csv"), PCB_MAX_ERRORS="50", **base), 29612)
    assert all(p.returncode == 0 for p in procs), outs
    rows = _csv_rows(tmp_path / "budget.csv")
    assert rows[0][0] == 16384 and rows[0][1] >= 50          # 1 dB: the first slice is enough
