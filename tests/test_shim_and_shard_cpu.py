"""CPU-only: the reference's own main() compiles against the shim header with the single edit
INTEGRATION.md describes; the conf.txt driver builds; the frame sharding + counter reduction
used for N>1 GPUs is exercised with world_size-2 gloo."""
import os
import re
import subprocess
import sys

import pytest

from conftest import ROOT

REF = "/root/reference/PolarCodes/PolarCodes.cpp"
HOST = os.path.join(ROOT, "polarcodes_b200", "host")


@pytest.fixture(scope="module")
def built():
    from polarcodes_b200 import build
    build.build()


@pytest.mark.ref
def test_reference_main_compiles_against_shim(built, tmp_path):
    if not os.path.exists(REF):
        pytest.skip("reference sources absent")
    src = open(REF).read()
    pat = r'#include "(Resources|Encoder|SCDecoder|SCDecoderMinsum|SCLDecoderMinsum)\.cpp"\s*\n'
    assert len(re.findall(pat, src)) == 5
    first = True

    def repl(m):
        nonlocal first
        if first:
            first = False
            return '#include <algorithm>\n#include <cfloat>\n#include <cmath>\n#include "polar_b200_shim.h"\n'
        return ""

    edited = re.sub(pat, repl, src)
    # everything else of the reference's translation unit is untouched
    assert edited.count("polar_b200_shim.h") == 1 and "myThread(testType" in edited
    cpp = tmp_path / "PolarCodes_shim.cpp"   # scratch copy, never enters the repo
    cpp.write_text(edited)
    exe = tmp_path / "PolarCodes_shim"
    lib = os.path.join(ROOT, "polarcodes_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-w", "-I", HOST, str(cpp), "-L", lib, "-lpolar_b200",
                           f"-Wl,-rpath,{lib}", "-Wl,-rpath,/usr/local/cuda/lib64", "-o", str(exe)])
    assert exe.exists()


def test_driver_builds_and_reports_missing_device(built, tmp_path):
    subprocess.check_call(["make", "-s", "-C", HOST])
    exe = os.path.join(HOST, "polarcodes_driver")
    (tmp_path / "conf.txt").write_text("1\n256\n128\n1000\n3\n2\n4\n8\n2\n2\n1.0 2.0\n")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True)
    # same prompts as PolarCodes.cpp:254-285
    assert "Input test type FER (1), throughput (2)" in r.stdout and "Input 2 SNRbs" in r.stdout
    import polarcodes_b200 as pb
    if pb.lib().pcb_device_count() == 0:
        assert r.returncode == 2 and "no CUDA device" in r.stderr   # loud, no CPU fallback


WORKER = r"""
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from polarcodes_b200.shard import shard_range, reduce_counters
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
r, w = dist.get_rank(), dist.get_world_size()
total = 1001
lo, hi = shard_range(total, r, w)
# stand-in for the device counters of this rank's frames: frame f is in error iff f % 7 == 0
fe = sum(1 for f in range(lo, hi) if f % 7 == 0)
be = sum(f % 5 for f in range(lo, hi))
tot = reduce_counters({"frames": hi - lo, "frame_errors": fe, "bit_errors": be})
assert tot["frames"] == total, tot
assert tot["frame_errors"] == sum(1 for f in range(total) if f % 7 == 0)
assert tot["bit_errors"] == sum(f % 5 for f in range(total))
dist.destroy_process_group()
print("ok", r)
"""


def test_shard_and_reduce_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script), ROOT], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("ok" in o for o in outs)


def test_shard_range_covers_everything():
    from polarcodes_b200.shard import shard_range
    for total in (0, 1, 7, 1000, 131072):
        for w in (1, 2, 3, 8):
            seen = []
            for r in range(w):
                lo, hi = shard_range(total, r, w)
                assert 0 <= lo <= hi <= total
                seen += list(range(lo, hi)) if total <= 1000 else [lo, hi]
            if total <= 1000:
                assert seen == list(range(total))
