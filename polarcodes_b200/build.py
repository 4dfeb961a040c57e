"""Builds polarcodes_b200/libpolar_b200.so (CUDA kernels + the C ABI) for sm_100a with nvcc.

In-tree build so the .so travels to the GPU box with the snapshot.  nvcc cross-compiles here
without a GPU.  Usage: python -m polarcodes_b200.build [--force]
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libpolar_b200.so")
OBJ = os.path.join(HERE, "build")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# -fmad=false: no mul+add contraction anywhere on the decode path (bit-exact LLR arithmetic);
# no --use_fast_math, denormals kept (ftz off is the default).
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off"]
CUDA_SOURCES = ["decode_generic.cu", "decode_scl8q.cu", "scl_select.cu", "decode_sc32.cu", "decode_sct.cu", "decode_sc8.cu", "aux_kernels.cu", "capi.cu"]
CXX_SOURCES = ["construct.cpp"]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "polar_b200.h"))
    headers.append(os.path.abspath(__file__))
    objs, cmds = [], []
    for src in CUDA_SOURCES + CXX_SOURCES:
        sp = os.path.join(CSRC, src)
        op = os.path.join(OBJ, src + ".o")
        objs.append(op)
        if force or _stale(op, [sp] + headers):
            cmds.append([_nvcc()] + ARCH + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", op])
    # the translation units are independent: compile them side by side (the two largest take about a minute each)
    jobs = max(1, min(len(cmds), int(os.environ.get("PCB_BUILD_JOBS", "0")) or (os.cpu_count() or 1)))
    with ThreadPoolExecutor(max_workers=jobs) as pool:
        for rc, cmd in zip(pool.map(lambda c: subprocess.call(c), cmds), cmds):
            if rc:
                raise subprocess.CalledProcessError(rc, cmd)
    if force or _stale(OUT, objs):
        subprocess.check_call([_nvcc()] + ARCH + ["-shared", "-o", OUT] + objs + ["-lcudart", "-ldl"])
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
