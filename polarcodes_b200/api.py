"""ctypes binding of include/polar_b200.h — the host-side mirror used by tests and bench.py.

The names follow the reference's own call sites (PolarCodes.cpp:119-188): `Code` is what the
loop builds once per SNR point (sigma, InfIndexes*, FrozenBits), `Decoder` wraps one
pcb_handle and exposes encode / decode / montecarlo.  Nothing here computes on the CPU except
the tiny per-SNR code construction, which is host code in the reference too; if the CUDA
library or a device is missing every compute call raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# PCB_LIB_PATH: another build of the same library (kernel experiments), never a different implementation
LIB_PATH = os.environ.get("PCB_LIB_PATH") or os.path.join(HERE, "libpolar_b200.so")

SC, SCMINSUM, SCLMINSUM = 1, 2, 3
STRICT, FP32, INT8 = 0, 1, 2


def int8_q(qfraction):
    """precision argument of the 8-bit min-sum mode with the scale 2 << qfraction (PCB_PREC_INT8_Q)."""
    return INT8 | (int(qfraction) << 8)

_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_fp = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_up = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = C.POINTER(C.c_uint64)

# every symbol include/polar_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "pcb_last_error": (C.c_char_p, []),
    "pcb_version": (C.c_char_p, []),
    "pcb_device_count": (C.c_int, []),
    "pcb_sigma": (C.c_double, [C.c_int, C.c_int, C.c_double]),
    "pcb_inf_indexes_awgn": (C.c_int, [C.c_int, C.c_int, C.c_double, _ip]),
    "pcb_inf_indexes_bec": (C.c_int, [C.c_int, C.c_int, _ip]),
    "pcb_frozen_mask": (C.c_int, [C.c_int, _ip, C.c_int, _ip]),
    "pcb_add_frozen_bits": (C.c_int, [_ip, _ip, C.c_int, _ip, C.c_int]),
    "pcb_crc8": (C.c_uint, [_ip, C.c_int]),
    "pcb_crc16": (C.c_uint, [_ip, C.c_int]),
    "pcb_append_crc": (C.c_int, [_ip, C.c_int, C.c_int]),
    "pcb_lane_ops_per_frame": (C.c_int, [_ip, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "pcb_alu_peak": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "pcb_create": (C.c_void_p, [C.c_int] * 7),
    "pcb_destroy": (None, [C.c_void_p]),
    "pcb_set_code": (C.c_int, [C.c_void_p, _ip, C.c_double]),
    "pcb_encode_batch": (C.c_int, [C.c_void_p, _ip, _ip, C.c_int, C.c_int]),
    "pcb_encode_batch_packed": (C.c_int, [C.c_void_p, _up, _up, C.c_int, C.c_int]),
    "pcb_encode_batch_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "pcb_decode_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "pcb_decode_batch_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "pcb_decode_batch_f32_packed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "pcb_decode_batch_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_float)]),
    "pcb_decode_batch_dev_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]),
    "pcb_montecarlo": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64, C.c_uint64, C.c_uint64, _u64p, _u64p, _u64p,
                                 C.POINTER(C.c_double)]),
    "pcb_generate_batch": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64, C.c_uint64, C.c_int, _fp, _ip]),
    "pcb_generate_batch_dev": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]),
    "pcb_count_errors_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, _u64p]),
    "pcb_comm_unique_id": (C.c_int, [C.c_void_p]),
    "pcb_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "pcb_comm_allreduce": (C.c_int, [C.c_void_p, _u64p, C.c_int]),
    "pcb_host_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int] + [C.POINTER(C.c_double)] * 3),
    "pcb_info": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_int)] * 5 + [_u64p]),
}

_lib = None


class PolarError(RuntimeError):
    pass


def lib():
    """Loads libpolar_b200.so (no fallback: a missing library is an error)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PolarError(f"{LIB_PATH} is not built: run `python -m polarcodes_b200.build` (needs nvcc)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise PolarError(f"polar_b200 error {rc}: {lib().pcb_last_error().decode()}")


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class Code:
    """sigma, information indexes and frozen mask exactly as PolarCodes.cpp:123-128 builds them."""

    def __init__(self, n, k, crc=0, snr_db=2.0, zcomputer=2):
        L = lib()
        self.n, self.k, self.crc, self.snr_db, self.zcomputer = n, k, crc, snr_db, zcomputer
        self.sigma = L.pcb_sigma(n, k, snr_db)
        if zcomputer == 1:   # InfIndexesBEC(n, k) ignores crc (PolarCodes.cpp:125)
            self.inf = np.zeros(k, np.int32)
            _check(L.pcb_inf_indexes_bec(n, k, self.inf))
        else:
            self.inf = np.zeros(k + crc, np.int32)
            _check(L.pcb_inf_indexes_awgn(n, k + crc, self.sigma, self.inf))
        self.mask = np.zeros(n, np.int32)
        _check(L.pcb_frozen_mask(n, self.inf, len(self.inf), self.mask))


def lane_ops_per_frame(mask, L, decoder):
    """(ALU lane-ops, LLR updates) of one frame under the SURVEY.md section 8(d) convention."""
    mask = _i32(mask)
    ops, upd = C.c_double(0), C.c_double(0)
    _check(lib().pcb_lane_ops_per_frame(mask, len(mask), L, decoder, C.byref(ops), C.byref(upd)))
    return ops.value, upd.value


def crc8(bits):
    b = _i32(bits)
    return lib().pcb_crc8(b, len(b))


def crc16(bits):
    b = _i32(bits)
    return lib().pcb_crc16(b, len(b))


def append_crc(xinf, k, crc):
    x = _i32(xinf).copy()
    _check(lib().pcb_append_crc(x, k, crc))
    return x


def add_frozen_bits(xinf, n, inf):
    xinf, inf = _i32(xinf), _i32(inf)
    x = np.zeros(n, np.int32)
    _check(lib().pcb_add_frozen_bits(xinf, x, n, inf, len(inf)))
    return x


class Decoder:
    """One pcb_handle: (GPU, n, k, crc, L, decoder).  decoder ids are the conf.txt values."""

    def __init__(self, n, k, crc=0, L=1, decoder=SCLMINSUM, precision=STRICT, device=0):
        self._h = None
        h = lib().pcb_create(device, n, k, crc, L, decoder, precision)
        if not h:
            raise PolarError(f"pcb_create failed: {lib().pcb_last_error().decode()}")
        self._h = h
        self.n, self.k, self.decoder = n, k, decoder
        self.crc = crc      # decoder 3 checks it; decoders 1, 2 only see it in the generated payload (PolarCodes.cpp:136-155)
        self.L = L if decoder == SCLMINSUM else 1
        self.stride = k + crc if decoder == SCLMINSUM else n
        self.words = (self.stride + 31) // 32

    def close(self):
        if self._h:
            lib().pcb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_code(self, mask, sigma):
        _check(lib().pcb_set_code(self._h, _i32(mask), float(sigma)))

    def encode(self, u, with_shuffle):
        u = _i32(u).reshape(-1, self.n)
        x = np.zeros_like(u)
        _check(lib().pcb_encode_batch(self._h, u, x, int(bool(with_shuffle)), u.shape[0]))
        return x

    def encode_packed(self, u_bits, with_shuffle):
        """u_bits [B][max(n/32,1)] uint32 -> x_bits, same shape (bit t of a frame at word t // 32, bit t % 32)."""
        u_bits = np.ascontiguousarray(u_bits, dtype=np.uint32).reshape(-1, max(self.n // 32, 1))
        x = np.zeros_like(u_bits)
        _check(lib().pcb_encode_batch_packed(self._h, u_bits, x, int(bool(with_shuffle)), u_bits.shape[0]))
        return x

    def encode_dev(self, u_bits_dev_ptr, x_bits_dev_ptr, with_shuffle, B):
        ms = C.c_float(0)
        _check(lib().pcb_encode_batch_dev(self._h, u_bits_dev_ptr, x_bits_dev_ptr, int(bool(with_shuffle)), B, C.byref(ms)))
        return ms.value

    def decode(self, y):
        """y [B][n] float64 (the reference's type) or float32 -> sol [B][stride] int32."""
        y = np.ascontiguousarray(y)
        if y.dtype != np.float32:
            y = np.ascontiguousarray(y, dtype=np.float64)
        y = y.reshape(-1, self.n)
        sol = np.zeros((y.shape[0], self.stride), np.int32)
        fn = lib().pcb_decode_batch_f32 if y.dtype == np.float32 else lib().pcb_decode_batch
        _check(fn(self._h, y.ctypes.data, sol.ctypes.data, y.shape[0]))
        return sol

    def decode_ptr(self, y_ptr, sol_ptr, B, f32=False):
        """Raw host pointers (e.g. pinned buffers): the reference-facing call bench.py times."""
        fn = lib().pcb_decode_batch_f32 if f32 else lib().pcb_decode_batch
        _check(fn(self._h, y_ptr, sol_ptr, B))

    def decode_packed(self, y):
        y = np.ascontiguousarray(y, dtype=np.float32).reshape(-1, self.n)
        sol = np.zeros((y.shape[0], self.words), np.uint32)
        _check(lib().pcb_decode_batch_f32_packed(self._h, y.ctypes.data, sol.ctypes.data, y.shape[0]))
        return sol

    def decode_dev(self, y_dev_ptr, y_is_f32, sol_bits_dev_ptr, B):
        """Device pointers; returns the kernel time in ms (CUDA events on the handle's stream)."""
        ms = C.c_float(0)
        _check(lib().pcb_decode_batch_dev(self._h, y_dev_ptr, int(y_is_f32), sol_bits_dev_ptr, B, C.byref(ms)))
        return ms.value

    def decode_dev_noalloc(self, y_dev_ptr, y_is_f32, sol_bits_dev_ptr, B):
        """pcb_decode_batch_dev without any allocation (static device scratch; see include/polar_b200.h)."""
        _check(lib().pcb_decode_batch_dev_noalloc(self._h, y_dev_ptr, int(y_is_f32), sol_bits_dev_ptr, B))

    def montecarlo(self, snr_db, frames, seed=0x5EED0000, first_frame_id=0):
        done, fe, be, sec = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0), C.c_double(0)
        _check(lib().pcb_montecarlo(self._h, float(snr_db), int(frames), int(seed), int(first_frame_id),
                                    C.byref(done), C.byref(fe), C.byref(be), C.byref(sec)))
        return dict(frames=done.value, frame_errors=fe.value, bit_errors=be.value, seconds=sec.value)

    def generate(self, snr_db, B, seed=0x5EED0000, first_frame_id=0):
        y = np.zeros((B, self.n), np.float32)
        tx = np.zeros((B, self.stride), np.int32)
        _check(lib().pcb_generate_batch(self._h, float(snr_db), int(seed), int(first_frame_id), B, y, tx))
        return y, tx

    def generate_dev(self, snr_db, B, y_dev_ptr, tx_dev_ptr=None, seed=0x5EED0000, first_frame_id=0):
        _check(lib().pcb_generate_batch_dev(self._h, float(snr_db), int(seed), int(first_frame_id), B, y_dev_ptr,
                                            tx_dev_ptr))

    def count_errors_dev(self, dec_bits_dev_ptr, tx_bits_dev_ptr, B, allreduce=False):
        """{frames, frame_errors, bit_errors} of a device-resident batch (summed over the ranks when allreduce)."""
        c = (C.c_uint64 * 3)(0, 0, 0)
        _check(lib().pcb_count_errors_dev(self._h, dec_bits_dev_ptr, tx_bits_dev_ptr, B, int(bool(allreduce)), c))
        return dict(frames=c[0], frame_errors=c[1], bit_errors=c[2])

    def comm_init(self, id128, rank, world):
        """Binds an NCCL communicator (one process per GPU); id128 = unique_id() of rank 0."""
        buf = (C.c_char * 128).from_buffer_copy(bytes(id128))
        _check(lib().pcb_comm_init(self._h, buf, rank, world))

    def comm_allreduce(self, counters):
        c = (C.c_uint64 * len(counters))(*[int(v) for v in counters])
        _check(lib().pcb_comm_allreduce(self._h, c, len(counters)))
        return [int(v) for v in c]

    def host_probe(self, y_host_ptr, frames):
        """Pinned H2D rate, host narrowing rate and the frames/s bound they give pcb_decode_batch."""
        a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
        _check(lib().pcb_host_probe(self._h, y_host_ptr, frames, C.byref(a), C.byref(b), C.byref(c)))
        return {"h2d_GBps": a.value, "narrow_Gvalues_per_s": b.value, "bound_frames_per_s": c.value}

    def alu_peak(self):
        v = C.c_double(0)
        _check(lib().pcb_alu_peak(self._h, C.byref(v)))
        return v.value

    def info(self):
        v = [C.c_int(0) for _ in range(5)]
        launches = C.c_uint64(0)
        _check(lib().pcb_info(self._h, *[C.byref(x) for x in v], C.byref(launches)))
        keys = ("sm_count", "warps_per_block", "blocks", "smem_per_block", "first_smem_layer")
        d = {k: x.value for k, x in zip(keys, v)}
        d["kernel_launches"] = launches.value
        return d


def comm_unique_id():
    """128-byte NCCL unique id (rank 0 creates it, every rank passes it to Decoder.comm_init)."""
    buf = (C.c_char * 128)()
    _check(lib().pcb_comm_unique_id(buf))
    return bytes(buf)


def unpack_bits(words, nbits):
    """[B][W] uint32 (bit t at word t/32, bit t%32) -> [B][nbits] int32."""
    w = np.ascontiguousarray(words, dtype=np.uint32)
    bits = np.unpackbits(w.view(np.uint8).reshape(w.shape[0], -1), axis=1, bitorder="little")
    return bits[:, :nbits].astype(np.int32)
