// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// Wrapper translation unit that compiles the UNMODIFIED reference sources (read in place
// from $(REF)/PolarCodes, i.e. /root/reference/PolarCodes; nothing is copied into this
// repo) into oracle/_ref/libpolar_ref.so and exposes the reference's own functions through
// a small C ABI so tests / golden-vector generation / the CPU-baseline timer can call them.
//
// The reference is one TU: PolarCodes.cpp #includes the other five .cpp files
// (PolarCodes.cpp:10-14).  `main` is renamed so the TU can live in a shared object.
// Build flags (oracle/Makefile): -std=c++17 -O2 -mavx2 -ffp-contract=off
//   -include algorithm -include cfloat -include cmath   (MSVC pulls those in transitively).
#define main polar_reference_main
#include "PolarCodes.cpp"
#undef main

#include <chrono>
#include <cstdint>
#include <cstring>

namespace {
struct RefBuffers {
    int n = 0, m = 0;
    int** encTree = nullptr;
    double** decTree1 = nullptr;
    int** decTree2 = nullptr;
};
RefBuffers g_buf;
int g_scl_n = 0, g_scl_L = 0, g_scl_k = 0, g_scl_crc = 0;

// Buffer shapes as in PolarCodes.cpp:103-117.
void ensure_buffers(int n) {
    if (g_buf.n == n) return;
    int m = 0;
    while ((1 << m) < n) m++;
    g_buf.n = n;
    g_buf.m = m;
    g_buf.encTree = new int*[m + 1];
    g_buf.decTree1 = new double*[m + 1];
    g_buf.decTree2 = new int*[m + 1];
    for (int i = 0; i <= m; i++) {
        int s = (n >> i) > 8 ? (n >> i) : 8;
        g_buf.encTree[i] = new int[s];
        g_buf.decTree1[i] = new double[s];
        g_buf.decTree2[i] = new int[s];
    }
    // PolarCodes.cpp:287-288
    reverseShuffleMem = new int[n];
    reverseShuffleMemDouble = new double[n];
}
}  // namespace

extern "C" {

// Resources.cpp:13-15 — the global mt19937 `eng` is reseeded for reproducibility.
void ref_seed(uint32_t s) { eng.seed(s); }

// PolarCodes.cpp:123
double ref_sigma(int n, int k, double snr_db) { return sqrt(n / (2. * k * pow(10, snr_db / 10))); }

// PolarCodes.cpp:59-74 / 77-92
void ref_inf_indexes_awgn(int n, int k, double sigma, int* out) {
    std::vector<int> v = InfIndexesAWGN(n, k, sigma);
    for (int i = 0; i < k; i++) out[i] = v[i];
}
void ref_inf_indexes_bec(int n, int k, int* out) {
    std::vector<int> v = InfIndexesBEC(n, k);
    for (int i = 0; i < k; i++) out[i] = v[i];
}

// Encoder.cpp:43-55 with an all-zero frozen vector (PolarCodes.cpp:127-128).
void ref_frozen_mask(int n, const int* inf, int ninf, int* mask) {
    std::vector<int> iv(inf, inf + ninf);
    std::vector<int> fz(n - ninf);
    std::vector<int> r = FrozenBits(n, iv, fz);
    for (int i = 0; i < n; i++) mask[i] = r[i];
}

// Encoder.cpp:17-33
void ref_add_frozen_bits(const int* xInf, int* x, int n, const int* inf, int ninf) {
    std::vector<int> iv(inf, inf + ninf);
    std::vector<int> fz(n - ninf);
    AddFrozenBits(const_cast<int*>(xInf), x, n, iv, fz);
}

// Encoder.cpp:77-83
void ref_encode(int n, const int* u, int* x, int with_shuffle) {
    ensure_buffers(n);
    Encode(n, const_cast<int*>(u), g_buf.encTree, with_shuffle != 0);
    std::memcpy(x, g_buf.encTree[0], sizeof(int) * n);
}

// Resources.cpp:108-138
unsigned ref_crc8(const int* data, int len) { return Crc8(const_cast<int*>(data), len); }
unsigned ref_crc16(const int* data, int len) { return Crc16(const_cast<int*>(data), len); }

// SCDecoder.cpp:74-85
void ref_sc_decode(int n, const double* y, const int* mask, double sigma, int* sol) {
    ensure_buffers(n);
    std::vector<int> fb(mask, mask + n);
    SCDecoder::Decode(n, const_cast<double*>(y), g_buf.decTree1, g_buf.decTree2, sol, fb, sigma);
}

// SCDecoderMinsum.cpp:68-76
void ref_scms_decode(int n, const double* y, const int* mask, int* sol) {
    ensure_buffers(n);
    std::vector<int> fb(mask, mask + n);
    SCDecoderMinsum::Decode(n, const_cast<double*>(y), g_buf.decTree1, g_buf.decTree2, sol, fb);
}

// SCLDecoderMinsum.cpp:60-118.  The namespace keeps global state; re-initialising for a new
// (n,L,k,crc) leaks the previous pools (as the reference would) but is otherwise sound once
// the path free-stack cursor is zeroed (it is a zero-initialised global the reference only
// ever increments in InitializeDataStructures, SCLDecoderMinsum.cpp:18,95).
void ref_scl_init(int n, int L, int k, int crc) {
    if (g_scl_n == n && g_scl_L == L && g_scl_k == k && g_scl_crc == crc) return;
    SCLDecoderMinsum::inactivePathIndicesEnd = 0;
    SCLDecoderMinsum::InitializeDataStructures(n, L, k, crc);
    g_scl_n = n; g_scl_L = L; g_scl_k = k; g_scl_crc = crc;
}

// SCLDecoderMinsum.cpp:401-477; sol has k+crc entries.
void ref_scl_decode(int n, int L, const double* y, const int* mask, int* sol) {
    ensure_buffers(n);
    std::vector<int> fb(mask, mask + n);
    SCLDecoderMinsum::Decode(n, L, const_cast<double*>(y), g_buf.decTree2, sol, fb, 2);
}

// B frames through the reference decoders, one call per frame exactly as the loop makes them
// (PolarCodes.cpp:162,171,180); sol is [B][n] for decoders 1,2 and [B][k+crc] for decoder 3.
void ref_decode_batch(int decoder, int n, int L, int k, int crc, double sigma, const double* y, const int* mask, int* sol, int B) {
    ensure_buffers(n);
    std::vector<int> fb(mask, mask + n);
    std::vector<int> tmp(n);
    if (decoder == 3) ref_scl_init(n, L, k, crc);
    const int stride = decoder == 3 ? k + crc : n;
    for (int f = 0; f < B; f++) {
        double* c = const_cast<double*>(y) + (size_t)f * n;
        switch (decoder) {
            case 1: SCDecoder::Decode(n, c, g_buf.decTree1, g_buf.decTree2, tmp.data(), fb, sigma); break;
            case 2: SCDecoderMinsum::Decode(n, c, g_buf.decTree1, g_buf.decTree2, tmp.data(), fb); break;
            default: SCLDecoderMinsum::Decode(n, L, c, g_buf.decTree2, tmp.data(), fb, 2); break;
        }
        std::memcpy(sol + (size_t)f * stride, tmp.data(), sizeof(int) * stride);
    }
}

// One iteration of the frame-generation part of the FER loop, PolarCodes.cpp:133-158:
// uniform info bits, CRC append (MSB first), AddFrozenBits, Encode, Transform (BPSK+AWGN).
// Draws from the global `eng` exactly as the reference does.
void ref_gen_frame(int n, int k, int crc, const int* inf, double sigma, int with_shuffle,
                   int* xInf /*k+crc*/, int* x /*n*/, int* cw /*n*/, double* y /*n*/) {
    ensure_buffers(n);
    std::uniform_int_distribution<int> distributionU(0, 1);
    std::normal_distribution<double> distributionN(0, sigma);
    for (int u = 0; u < k; u++) xInf[u] = distributionU(eng);
    if (crc == 8) {
        unsigned short crcRes = Crc8(xInf, k);
        for (int u = 1; u <= 8; u++) { xInf[k + crc - u] = crcRes & 0x01; crcRes >>= 1; }
    } else if (crc == 16) {
        unsigned short crcRes = Crc16(xInf, k);
        for (int u = 1; u <= 16; u++) { xInf[k + crc - u] = crcRes & 0x0001; crcRes >>= 1; }
    }
    std::vector<int> iv(inf, inf + k + crc);
    std::vector<int> fz(n - k - crc);
    AddFrozenBits(xInf, x, n, iv, fz);
    Encode(n, x, g_buf.encTree, with_shuffle != 0);
    std::memcpy(cw, g_buf.encTree[0], sizeof(int) * n);
    Transform(n, g_buf.encTree[0], y, distributionN);
}

// Decode-only timing over a batch of DISTINCT frames (the "fair" CPU baseline of
// SURVEY.md §8(d)); returns wall seconds (steady_clock).  decoder: 1 SC, 2 SCMinsum, 3 SCL.
double ref_time_decode(int decoder, int n, int L, int k, int crc, const double* y, const int* mask,
                       double sigma, int frames, int reps, int* sol_last) {
    ensure_buffers(n);
    std::vector<int> fb(mask, mask + n);
    std::vector<int> sol(n);
    if (decoder == 3) ref_scl_init(n, L, k, crc);
    auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < reps; r++) {
        for (int f = 0; f < frames; f++) {
            double* c = const_cast<double*>(y) + (size_t)f * n;
            switch (decoder) {
                case 1: SCDecoder::Decode(n, c, g_buf.decTree1, g_buf.decTree2, sol.data(), fb, sigma); break;
                case 2: SCDecoderMinsum::Decode(n, c, g_buf.decTree1, g_buf.decTree2, sol.data(), fb); break;
                default: SCLDecoderMinsum::Decode(n, L, c, g_buf.decTree2, sol.data(), fb, 2); break;
            }
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    if (sol_last) std::memcpy(sol_last, sol.data(), sizeof(int) * (decoder == 3 ? k + crc : n));
    return std::chrono::duration<double>(t1 - t0).count();
}

// The reference's whole FER loop for one SNR point (PolarCodes.cpp:122-195) with the global
// engine; returns the frame-error count.  zcomputer: 1 BEC, 2 AWGN.
int ref_fer_point(int n, int k, int crc, int L, int decoder, int zcomputer, double snr_db, int loop) {
    ensure_buffers(n);
    double sigma = ref_sigma(n, k, snr_db);
    std::vector<int> inf = zcomputer == 1 ? InfIndexesBEC(n, k) : InfIndexesAWGN(n, k + crc, sigma);
    std::vector<int> fz(n - k - crc);
    std::vector<int> mask = FrozenBits(n, inf, fz);
    if (decoder == 3) ref_scl_init(n, L, k, crc);
    std::vector<int> xInf(k + crc), x(n), cw(n), dec(n);
    std::vector<double> y(n);
    int er = 0;
    for (int i = 0; i < loop; i++) {
        ref_gen_frame(n, k, crc, inf.data(), sigma, decoder != 3, xInf.data(), x.data(), cw.data(), y.data());
        if (decoder == 1) {
            SCDecoder::Decode(n, y.data(), g_buf.decTree1, g_buf.decTree2, dec.data(), mask, sigma);
            for (int u = 0; u < n; u++) if (x[u] != dec[u]) { er++; break; }
        } else if (decoder == 2) {
            SCDecoderMinsum::Decode(n, y.data(), g_buf.decTree1, g_buf.decTree2, dec.data(), mask);
            for (int u = 0; u < n; u++) if (x[u] != dec[u]) { er++; break; }
        } else {
            SCLDecoderMinsum::Decode(n, L, y.data(), g_buf.decTree2, dec.data(), mask, 2);
            for (int u = 0; u < k; u++) if (xInf[u] != dec[u]) { er++; break; }
        }
    }
    return er;
}

// the reference's 8-bit sketches (SCDecoderMinsum.cpp:17-30), never called by its own decoders
int ref_decodef_char(int r1, int r2) { return (int)SCDecoderMinsum::Decodef_char((char)r1, (char)r2); }
int ref_decodeg_char(int r1, int r2, int b) { return (int)(signed char)SCDecoderMinsum::Decodeg_char((char)r1, (char)r2, b); }
}  // extern "C"
