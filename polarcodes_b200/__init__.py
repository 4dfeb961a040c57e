"""polarcodes_b200 — B200-native batched polar encoder / SC / SCL decoder behind the
dim49v/polarCodes function interfaces.  CUDA kernels + C ABI live in csrc/ (built into
libpolar_b200.so by polarcodes_b200.build); api.py is the ctypes mirror used by the tests
and bench.py; host/ holds the C++ shim and the conf.txt driver."""
from .api import (FP32, INT8, SC, SCLMINSUM, SCMINSUM, STRICT, Code, Decoder, PolarError, add_frozen_bits,  # noqa: F401
                  append_crc, comm_unique_id, crc8, crc16, int8_q, lane_ops_per_frame, lib, unpack_bits)
