// polar_b200_shim.h — the reference's own function signatures on top of the C ABI.
//
// dim49v/polarCodes has no plugin layer: PolarCodes.cpp #includes Encoder.cpp, SCDecoder.cpp,
// SCDecoderMinsum.cpp, SCLDecoderMinsum.cpp and Resources.cpp (PolarCodes.cpp:10-14) and calls
// their free functions.  Replacing those five #include lines by this header (the only edit,
// see INTEGRATION.md) makes the unchanged Monte-Carlo loop run every encode/decode on the GPU,
// one frame per call (B = 1).  That is for drop-in correctness; throughput comes from the
// batched entry points (pcb_decode_batch / pcb_montecarlo, host/polarcodes_driver.cpp).
//
// Signatures, argument meaning and buffer ownership are the reference's (file:line cited per
// function).  Scratch arguments the GPU path does not need (encTree, decTree1, decTree2) are
// accepted and, where the reference returns its result through them, filled in.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>
#include <tuple>
#include <vector>

#include "../../include/polar_b200.h"

// Resources.cpp:13-17 — the global engine and the ReverseShuffle scratch the driver allocates
static std::random_device polar_shim_rd;
static std::mt19937 eng(polar_shim_rd());
static int* reverseShuffleMem = nullptr;
static double* reverseShuffleMemDouble = nullptr;

namespace polar_shim {

inline void die(const char* where) {
    std::fprintf(stderr, "polar_b200 shim: %s failed: %s\n", where, pcb_last_error());
    std::exit(2);   // the reference has no error channel; a GPU failure must not pass silently
}

// one handle per (n, k, crc, L, decoder); the frozen mask / sigma are re-uploaded when they change
struct Slot {
    pcb_handle* h = nullptr;
    std::vector<int> mask;
    double sigma = -1;
};

inline Slot& slot(int n, int k, int crc, int L, int decoder) {
    static std::map<std::tuple<int, int, int, int, int>, Slot> slots;
    Slot& s = slots[std::make_tuple(n, k, crc, L, decoder)];
    if (!s.h) {
        s.h = pcb_create(0, n, k, crc, L, decoder, PCB_PREC_STRICT);
        if (!s.h) die("pcb_create");
    }
    return s;
}

inline void set_code(Slot& s, const std::vector<int>& frozenBits, double sigma) {
    if (s.mask != frozenBits || s.sigma != sigma) {
        if (pcb_set_code(s.h, frozenBits.data(), sigma) != PCB_OK) die("pcb_set_code");
        s.mask = frozenBits;
        s.sigma = sigma;
    }
}

inline int count_info(const std::vector<int>& frozenBits) {
    int c = 0;
    for (int v : frozenBits) c += (v == -1);
    return c;
}

}  // namespace polar_shim

// ---- Resources.cpp -------------------------------------------------------------------
inline unsigned char Crc8(int* data, int len) { return (unsigned char)pcb_crc8(data, len); }       // :108-122
inline unsigned short Crc16(int* data, int len) { return (unsigned short)pcb_crc16(data, len); }   // :124-138

// ---- Encoder.cpp ---------------------------------------------------------------------
// :17-33
inline void AddFrozenBits(int* xInf, int* x, int n, const std::vector<int>& infIndexes, const std::vector<int>& frozenBits) {
    (void)frozenBits;   // all zero in the reference (PolarCodes.cpp:127)
    if (pcb_add_frozen_bits(xInf, x, n, infIndexes.data(), (int)infIndexes.size()) != PCB_OK) polar_shim::die("pcb_add_frozen_bits");
}
// :43-55
inline std::vector<int> FrozenBits(int n, const std::vector<int>& infIndexes, const std::vector<int>& frozenBits) {
    (void)frozenBits;
    std::vector<int> mask(n);
    if (pcb_frozen_mask(n, infIndexes.data(), (int)infIndexes.size(), mask.data()) != PCB_OK) polar_shim::die("pcb_frozen_mask");
    return mask;
}
// :35-41 — host RNG exactly as the reference (BPSK 0 -> +1, 1 -> -1, plus N(0, sigma))
inline void Transform(int n, int* c, double* res, std::normal_distribution<double>& distribution) {
    for (int i = 0; i < n; i++) res[i] = (c[i] == 0 ? 1. : -1.) + distribution(eng);
}
// :77-83 — result is left in encTree[0][0..n)
inline void Encode(int n, int* c, int** encTree, bool withShuffle) {
    polar_shim::Slot& s = polar_shim::slot(n, 1, 0, 1, PCB_DEC_SCMINSUM);
    if (pcb_encode_batch(s.h, c, encTree[0], withShuffle ? 1 : 0, 1) != PCB_OK) polar_shim::die("pcb_encode_batch");
}

// ---- decoders --------------------------------------------------------------------------
namespace SCDecoder {
// SCDecoder.cpp:74-85
inline void Decode(int n, double* c, double** decTree1, int** decTree2, int* sol, const std::vector<int>& frozenBits, double sigma) {
    (void)decTree1; (void)decTree2;
    polar_shim::Slot& s = polar_shim::slot(n, polar_shim::count_info(frozenBits), 0, 1, PCB_DEC_SC);
    polar_shim::set_code(s, frozenBits, sigma);
    if (pcb_decode_batch(s.h, c, sol, 1) != PCB_OK) polar_shim::die("pcb_decode_batch");
}
}  // namespace SCDecoder

namespace SCDecoderMinsum {
// SCDecoderMinsum.cpp:68-76
inline void Decode(int n, double* c, double** decTree1, int** decTree2, int* sol, const std::vector<int>& frozenBits) {
    (void)decTree1; (void)decTree2;
    polar_shim::Slot& s = polar_shim::slot(n, polar_shim::count_info(frozenBits), 0, 1, PCB_DEC_SCMINSUM);
    polar_shim::set_code(s, frozenBits, 1.0);
    if (pcb_decode_batch(s.h, c, sol, 1) != PCB_OK) polar_shim::die("pcb_decode_batch");
}
}  // namespace SCDecoderMinsum

namespace SCLDecoderMinsum {
static int shim_n = 0, shim_L = 0, shim_k = 0, shim_crc = 0;
// SCLDecoderMinsum.cpp:60-118 — remembers (n, L, k, crc); device state is created lazily
inline void InitializeDataStructures(int n, int L, int k, int crc) { shim_n = n; shim_L = L; shim_k = k; shim_crc = crc; }
// SCLDecoderMinsum.cpp:401-477 — writes sol[0..k+crc)
inline void Decode(int n, int L, double* c, int** decTree2, int* sol, const std::vector<int>& frozenBits, int qFraction) {
    (void)decTree2; (void)qFraction;   // qFraction is dead in the reference too (:405)
    polar_shim::Slot& s = polar_shim::slot(n, shim_k, shim_crc, L, PCB_DEC_SCLMINSUM);
    polar_shim::set_code(s, frozenBits, 1.0);
    if (pcb_decode_batch(s.h, c, sol, 1) != PCB_OK) polar_shim::die("pcb_decode_batch");
}
}  // namespace SCLDecoderMinsum
