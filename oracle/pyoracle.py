"""TEST INFRASTRUCTURE ONLY — ctypes bindings for the CPU oracle.

`Oracle`  wraps oracle/liboracle.so  (plain-C restatement, polar_oracle.c; kind "port").
`Ref`     wraps oracle/_ref/libpolar_ref.so (the unmodified reference compiled in place by
          oracle/Makefile; kind "reference").  Both expose the same methods so a test can
          run either against the CUDA path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpolar_ref.so")

_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


SHIM_MAIN = os.path.join(HERE, "_ref", "PolarCodes_shim")


def build(force=False):
    """Compile liboracle.so (always possible) and, only where /root/reference exists, oracle/_ref:
    libpolar_ref.so (the unmodified reference) and PolarCodes_shim (its main() against the shim header)."""
    def stale(out, src):
        return not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src)
    need = force or stale(ORACLE_SO, os.path.join(HERE, "polar_oracle.c"))
    if os.path.isdir("/root/reference/PolarCodes"):
        shim_h = os.path.join(HERE, "..", "polarcodes_b200", "host", "polar_b200_shim.h")
        need = need or stale(REF_SO, os.path.join(HERE, "ref_wrap.cpp")) or stale(SHIM_MAIN, shim_h)
    if need:
        subprocess.check_call(["make", "-C", HERE, "all"], stdout=subprocess.DEVNULL)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class _Base:
    kind = "?"

    def make_code(self, n, k, crc, snr_db, zcomputer=2):
        """(sigma, info indexes, frozen mask) exactly as PolarCodes.cpp:123-128."""
        sigma = self.sigma(n, k, snr_db)
        inf = self.inf_indexes_bec(n, k) if zcomputer == 1 else self.inf_indexes_awgn(n, k + crc, sigma)
        return sigma, inf, self.frozen_mask(n, inf)


class Oracle(_Base):
    kind = "port"

    def __init__(self):
        build()
        L = self.lib = C.CDLL(ORACLE_SO)
        L.po_sigma.restype = C.c_double
        L.po_sigma.argtypes = [C.c_int, C.c_int, C.c_double]
        L.po_inf_indexes_awgn.argtypes = [C.c_int, C.c_int, C.c_double, _ip]
        L.po_inf_indexes_bec.argtypes = [C.c_int, C.c_int, _ip]
        L.po_frozen_mask.argtypes = [C.c_int, _ip, C.c_int, _ip]
        L.po_add_frozen_bits.argtypes = [_ip, _ip, C.c_int, _ip, C.c_int]
        L.po_encode.argtypes = [C.c_int, _ip, _ip, C.c_int]
        L.po_crc8.restype = C.c_uint
        L.po_crc8.argtypes = [_ip, C.c_int]
        L.po_crc16.restype = C.c_uint
        L.po_crc16.argtypes = [_ip, C.c_int]
        L.po_append_crc.argtypes = [_ip, C.c_int, C.c_int]
        L.po_decode_batch.restype = C.c_int
        L.po_decode_batch.argtypes = [C.c_int] * 5 + [C.c_double, _dp, _ip, _ip, C.c_int]
        L.po_time_decode.restype = C.c_double
        L.po_time_decode.argtypes = [C.c_int] * 5 + [C.c_double, _dp, _ip, C.c_int, C.c_int]

    def sigma(self, n, k, snr_db):
        return self.lib.po_sigma(n, k, snr_db)

    def inf_indexes_awgn(self, n, k, sigma):
        out = np.zeros(k, np.int32)
        self.lib.po_inf_indexes_awgn(n, k, sigma, out)
        return out

    def inf_indexes_bec(self, n, k):
        out = np.zeros(k, np.int32)
        self.lib.po_inf_indexes_bec(n, k, out)
        return out

    def frozen_mask(self, n, inf):
        inf = _i32(inf)
        out = np.zeros(n, np.int32)
        self.lib.po_frozen_mask(n, inf, len(inf), out)
        return out

    def add_frozen_bits(self, xinf, n, inf):
        xinf, inf = _i32(xinf), _i32(inf)
        x = np.zeros(n, np.int32)
        self.lib.po_add_frozen_bits(xinf, x, n, inf, len(inf))
        return x

    def encode(self, u, with_shuffle):
        u = _i32(u)
        x = np.zeros_like(u)
        for f in range(u.reshape(-1, u.shape[-1]).shape[0]):
            self.lib.po_encode(u.shape[-1], u.reshape(-1, u.shape[-1])[f], x.reshape(-1, u.shape[-1])[f],
                               int(with_shuffle))
        return x

    def crc8(self, bits, length=None):
        bits = _i32(bits)
        return self.lib.po_crc8(bits, len(bits) if length is None else length)

    def crc16(self, bits, length=None):
        bits = _i32(bits)
        return self.lib.po_crc16(bits, len(bits) if length is None else length)

    def append_crc(self, xinf, k, crc):
        xinf = _i32(xinf).copy()
        self.lib.po_append_crc(xinf, k, crc)
        return xinf

    def decode(self, decoder, n, L, k, crc, sigma, y, mask):
        y = _f64(y).reshape(-1, n)
        stride = k + crc if decoder == 3 else n
        sol = np.zeros((y.shape[0], stride), np.int32)
        rc = self.lib.po_decode_batch(decoder, n, L, k, crc, sigma, y, _i32(mask), sol, y.shape[0])
        assert rc == 0
        return sol

    def decode_int8(self, n, qfraction, y, mask):
        """8-bit min-sum SC (polar_oracle.c: the definition of the product's PCB_PREC_INT8 mode) -> [B][n] decisions."""
        y = _f64(y).reshape(-1, n)
        sol = np.zeros((y.shape[0], n), np.int32)
        self.lib.po_sc8_decode_batch.argtypes = [C.c_int, C.c_int, _dp, _ip, _ip, C.c_int]
        assert self.lib.po_sc8_decode_batch(n, qfraction, y, _i32(mask), sol, y.shape[0]) == 0
        return sol

    def time_decode(self, decoder, n, L, k, crc, sigma, y, mask, reps=1):
        y = _f64(y).reshape(-1, n)
        return self.lib.po_time_decode(decoder, n, L, k, crc, sigma, y, _i32(mask), y.shape[0], reps)


class Ref(_Base):
    kind = "reference"

    @staticmethod
    def available():
        if not os.path.exists(REF_SO) and os.path.isdir("/root/reference/PolarCodes"):
            build()
        return os.path.exists(REF_SO)

    def __init__(self):
        if not Ref.available():
            raise RuntimeError("oracle/_ref/libpolar_ref.so not built (reference sources absent)")
        L = self.lib = C.CDLL(REF_SO)
        L.ref_seed.argtypes = [C.c_uint32]
        L.ref_sigma.restype = C.c_double
        L.ref_sigma.argtypes = [C.c_int, C.c_int, C.c_double]
        L.ref_inf_indexes_awgn.argtypes = [C.c_int, C.c_int, C.c_double, _ip]
        L.ref_inf_indexes_bec.argtypes = [C.c_int, C.c_int, _ip]
        L.ref_frozen_mask.argtypes = [C.c_int, _ip, C.c_int, _ip]
        L.ref_add_frozen_bits.argtypes = [_ip, _ip, C.c_int, _ip, C.c_int]
        L.ref_encode.argtypes = [C.c_int, _ip, _ip, C.c_int]
        L.ref_crc8.restype = C.c_uint
        L.ref_crc8.argtypes = [_ip, C.c_int]
        L.ref_crc16.restype = C.c_uint
        L.ref_crc16.argtypes = [_ip, C.c_int]
        L.ref_sc_decode.argtypes = [C.c_int, _dp, _ip, C.c_double, _ip]
        L.ref_scms_decode.argtypes = [C.c_int, _dp, _ip, _ip]
        L.ref_scl_init.argtypes = [C.c_int] * 4
        L.ref_scl_decode.argtypes = [C.c_int, C.c_int, _dp, _ip, _ip]
        L.ref_gen_frame.argtypes = [C.c_int, C.c_int, C.c_int, _ip, C.c_double, C.c_int, _ip, _ip, _ip, _dp]
        L.ref_time_decode.restype = C.c_double
        L.ref_time_decode.argtypes = [C.c_int] * 5 + [_dp, _ip, C.c_double, C.c_int, C.c_int, C.c_void_p]
        L.ref_decode_batch.argtypes = [C.c_int] * 5 + [C.c_double, _dp, _ip, _ip, C.c_int]
        L.ref_fer_point.restype = C.c_int
        L.ref_fer_point.argtypes = [C.c_int] * 6 + [C.c_double, C.c_int]

    def seed(self, s):
        self.lib.ref_seed(s)

    def sigma(self, n, k, snr_db):
        return self.lib.ref_sigma(n, k, snr_db)

    def inf_indexes_awgn(self, n, k, sigma):
        out = np.zeros(k, np.int32)
        self.lib.ref_inf_indexes_awgn(n, k, sigma, out)
        return out

    def inf_indexes_bec(self, n, k):
        out = np.zeros(k, np.int32)
        self.lib.ref_inf_indexes_bec(n, k, out)
        return out

    def frozen_mask(self, n, inf):
        inf = _i32(inf)
        out = np.zeros(n, np.int32)
        self.lib.ref_frozen_mask(n, inf, len(inf), out)
        return out

    def add_frozen_bits(self, xinf, n, inf):
        xinf, inf = _i32(xinf), _i32(inf)
        x = np.zeros(n, np.int32)
        self.lib.ref_add_frozen_bits(xinf, x, n, inf, len(inf))
        return x

    def encode(self, u, with_shuffle):
        u = _i32(u)
        n = u.shape[-1]
        x = np.zeros_like(u)
        uu, xx = u.reshape(-1, n), x.reshape(-1, n)
        for f in range(uu.shape[0]):
            self.lib.ref_encode(n, uu[f], xx[f], int(with_shuffle))
        return x

    def crc8(self, bits, length=None):
        bits = _i32(bits)
        return self.lib.ref_crc8(bits, len(bits) if length is None else length)

    def crc16(self, bits, length=None):
        bits = _i32(bits)
        return self.lib.ref_crc16(bits, len(bits) if length is None else length)

    def decode(self, decoder, n, L, k, crc, sigma, y, mask):
        y = _f64(y).reshape(-1, n)
        stride = k + crc if decoder == 3 else n
        sol = np.zeros((y.shape[0], stride), np.int32)
        self.lib.ref_decode_batch(decoder, n, L, k, crc, float(sigma), y, _i32(mask), sol, y.shape[0])
        return sol

    def gen_frames(self, n, k, crc, inf, sigma, with_shuffle, frames):
        """frames iterations of PolarCodes.cpp:133-158 drawing from the (seeded) global engine."""
        inf = _i32(inf)
        xinf = np.zeros((frames, k + crc), np.int32)
        x = np.zeros((frames, n), np.int32)
        cw = np.zeros((frames, n), np.int32)
        y = np.zeros((frames, n), np.float64)
        for f in range(frames):
            self.lib.ref_gen_frame(n, k, crc, inf, sigma, int(with_shuffle), xinf[f], x[f], cw[f], y[f])
        return xinf, x, cw, y

    def time_decode(self, decoder, n, L, k, crc, sigma, y, mask, reps=1):
        y = _f64(y).reshape(-1, n)
        return self.lib.ref_time_decode(decoder, n, L, k, crc, y, _i32(mask), sigma, y.shape[0], reps, None)

    def fer_point(self, n, k, crc, L, decoder, zcomputer, snr_db, loop):
        return self.lib.ref_fer_point(n, k, crc, L, decoder, zcomputer, snr_db, loop)


def _ref_worker(args):
    decoder, n, L, k, crc, sigma, y, mask = args
    return Ref().decode(decoder, n, L, k, crc, sigma, y, mask)


def ref_decode_parallel(decoder, n, L, k, crc, sigma, y, mask, procs=None):
    """The unmodified reference over a large batch, split over host processes (its list decoder keeps
    namespace-global state: processes, not threads).  Frame order is preserved."""
    import multiprocessing as mp
    y = _f64(y).reshape(-1, n)
    procs = max(1, min(procs or (os.cpu_count() or 1), y.shape[0] // 8 or 1))
    if procs == 1:
        return Ref().decode(decoder, n, L, k, crc, sigma, y, mask)
    parts = np.array_split(y, procs)
    with mp.get_context("fork").Pool(procs) as pool:
        out = pool.map(_ref_worker, [(decoder, n, L, k, crc, sigma, p, mask) for p in parts])
    return np.concatenate(out, axis=0)


def best_cpu():
    """The strongest available checker: the compiled reference if present, else the C port."""
    return Ref() if Ref.available() else Oracle()
