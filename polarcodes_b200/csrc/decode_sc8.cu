// 8-bit min-sum SC decoder, four LLRs per 32-bit register (PCB_PREC_INT8).
//
// The reference sketches this mode and never wires it up: Decodef_char / Decodeg_char (SCDecoderMinsum.cpp:17-30) and
// the quantiser scale `fraction = 2 << qFraction` (SCLDecoderMinsum.cpp:405, conf.txt:9).  The definition implemented
// here is po_sc8_decode_batch of the CPU restatement under oracle/ (checked against the reference's two char functions): the tree
// walk of SCDecoderMinsum.cpp:32-66 on int8 LLRs, q(y) = clamp(rint(y * (2 << qFraction)), -127, 127), f = Decodef_char,
// g = Decodeg_char with the sum clamped to +-127 instead of wrapping.  It is a mode of its own -- not bit-exact with the
// double decoder -- and bit-exact with that definition.
//
// Mapping: lane = frame, a tile of 32 frames per one-warp block, as decode_sct.cu; a register holds FOUR consecutive
// elements of a level (one byte each), so a level of S elements is S/4 rows of 32 lanes and the pairing (x, x + S) of the
// half-split order pairs whole registers.  f and g are SWAR / packed-16-bit arithmetic on those registers.  The whole
// level store of a tile (n - 8 bytes per frame) and the heap live in shared memory; the channel values are quantised,
// bit-reversed and tile-transposed once per batch by sc8_quantise_kernel.  The tree is pruned at all-frozen subtrees
// (exact in any arithmetic); the all-information shortcut of decode_sct.cu is NOT used (8-bit LLRs are zero too often).
#include <cstdio>

#include "decode_common.cuh"
#include "decode_launch.h"
#include "sc_ops.cuh"
#include "sct_common.cuh"

#include <cstdlib>

#ifndef PCB_8_QU
#define PCB_8_QU 4           // frames per iteration of the fused quantiser (4 x 4 loads in flight per lane)
#endif
#ifndef PCB_8_VS_DEFAULT
#define PCB_8_VS_DEFAULT 5   // the first level kept in shared memory is at most level m - 5 (64 values)
#endif

namespace pcb {
namespace sc8 {

// every byte -> 0xFF if its sign bit is set, else 0
// prmt with the full selector nibbles (bit 3 = replicate the sign of the selected byte; __byte_perm keeps only 3 bits)
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
__device__ __forceinline__ uint32_t smask(uint32_t a) { return prmt(a, 0u, 0xba98u); }
// per byte: x negated where m = 0xFF (x in [-127, 127]; the add cannot carry out of a byte)
__device__ __forceinline__ uint32_t negm(uint32_t x, uint32_t m) {
    const uint32_t t = x ^ m;
    return (((t & 0x7f7f7f7fu) + (m & 0x01010101u)) ^ (t & 0x80808080u));
}
// f on four bytes: min(|a|, |b|), negative iff the sign bits differ (Decodef_char)
__device__ __forceinline__ uint32_t f4(uint32_t a, uint32_t b) {
    const uint32_t sa = smask(a), sb = smask(b);
    const uint32_t aa = (a ^ sa) + (sa & 0x01010101u), ab = (b ^ sb) + (sb & 0x01010101u);      // |.| <= 127: no carries
    const uint32_t ge = smask((aa | 0x80808080u) - ab);                                          // 0xFF where |a| >= |b| (no borrows)
    const uint32_t mn = (ab & ge) | (aa & ~ge);
    return negm(mn, sa ^ sb);
}
// g on four bytes: clamp(b + (c ? -a : a), -127, 127); nib = the four partial-sum bits
__device__ __forceinline__ uint32_t g4(uint32_t a, uint32_t b, uint32_t nib) {
    const uint32_t m = ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;                              // bit i -> byte i
    const uint32_t an = negm(a, m);
    // the sums in 16 bits (two per register), clamped, packed back
    const uint32_t alo = prmt(an, 0u, 0x9180u), ahi = prmt(an, 0u, 0xb3a2u);
    const uint32_t blo = prmt(b, 0u, 0x9180u), bhi = prmt(b, 0u, 0xb3a2u);
    const uint32_t slo = __vmaxs2(__vmins2(__vadd2(alo, blo), 0x007f007fu), 0xff81ff81u);
    const uint32_t shi = __vmaxs2(__vmins2(__vadd2(ahi, bhi), 0x007f007fu), 0xff81ff81u);
    return prmt(slo, shi, 0x6420u);
}

struct Args {
    DecodeArgs a;
    const uint32_t* ops;
    int nops;
    uint32_t* yq;            // quantised channel values, bit-reversed order, four per word: [tile][n/4][32] from the pre-pass, or
                             // (fused) one [n/4][32] slice per block that the tile's prologue fills
    int fused;               // the decoding warp quantises its own tile (no pre-pass)
    float scale;
    uint32_t* glev;          // per block: levels 1 .. vs - 1 (stay in L2), which shrinks the shared memory of a tile
    int vs;                  // first level in shared memory (2 or more)
    uint32_t* bits;          // [B][n/32] packed decisions
    int ntiles;
};

// q(y) = clamp(rint(y * scale), -127, 127) in double as the definition has it (oracle po_q8: round to nearest even, NaN -> -127):
// the conversion instruction rounds the same way and saturates, the compare sends NaN and everything below -127 to -127
__device__ __forceinline__ int q8i(double v, double sc) {
    const double p = v * sc;
    return p >= -127. ? min(127, __double2int_rn(p)) : -127;
}
__device__ __forceinline__ uint32_t q8(double v, double sc) { return (uint32_t)q8i(v, sc) & 0xFFu; }

// y[B][n] (double or float) -> yq[tile][x4][lane]: byte j of the word = q(y[frame][bitrev(4 x4 + j)]), frame = tile * 32 + lane
template <typename YT>
__global__ void __launch_bounds__(256) sc8_quantise_kernel(const YT* __restrict__ y, uint32_t* __restrict__ yq, int n, int m, int B, int ntiles, float scale) {
    extern __shared__ uint32_t sm[];                   // [32 frames][n/4 + 1] words
    const int W4 = n >> 2, stride = W4 + 1;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll 2
        for (int idx = threadIdx.x; idx < 32 * W4; idx += 256) {      // four consecutive source values per thread: coalesced
            const int f = idx / W4, c = idx - f * W4;
            const int frame = min(tile * 32 + f, B - 1);
            const YT* src = y + (size_t)frame * n + 4 * c;
            int v[4];
            if constexpr (sizeof(YT) == 8) {
                const double2 p0 = *reinterpret_cast<const double2*>(src), p1 = *reinterpret_cast<const double2*>(src + 2);
                v[0] = q8i(p0.x, (double)scale); v[1] = q8i(p0.y, (double)scale); v[2] = q8i(p1.x, (double)scale); v[3] = q8i(p1.y, (double)scale);
            } else {
                const float4 p = *reinterpret_cast<const float4*>(src);
                v[0] = q8i((double)p.x, (double)scale); v[1] = q8i((double)p.y, (double)scale); v[2] = q8i((double)p.z, (double)scale); v[3] = q8i((double)p.w, (double)scale);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int x = (int)(__brev((unsigned)(4 * c + j)) >> (32 - m));
                reinterpret_cast<signed char*>(sm + f * stride)[x] = (signed char)v[j];
            }
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < 32 * W4; idx += 256) {
            const int x4 = idx >> 5, lane = idx & 31;
            yq[((size_t)tile * W4 + x4) * 32 + lane] = sm[lane * stride + x4];
        }
        __syncthreads();
    }
}

// The quantiser as the prologue of a tile (same mapping as sc8_quantise_kernel): lane l of step g reads the values at offset
// o = 32 g + l of the four quarters of a frame (coalesced), packs them into word x4 = bitrev(o) of that frame (byte j holds
// quarter bitrev2(j)), and a 32 x 32 word transpose through shared memory turns 32 frames into rows [x4][lane = frame].
// The DRAM reads of one tile then overlap the decoding of the other tiles of the SM instead of being a pass of their own.
template <typename YT>
__device__ __forceinline__ void quantise_tile(const YT* __restrict__ y, int n, int m, int B, int tile, double sc, uint32_t* __restrict__ slice,
                                              uint32_t* tb, int lane) {
    const int Q = n >> 2;
    constexpr int QU = PCB_8_QU;
    for (int g0 = 0; g0 < Q; g0 += 32) {
        const int o = g0 + lane;
        const bool act = o < Q;
#pragma unroll QU
        for (int f = 0; f < 32; f++) {
            const int frame = min(tile * 32 + f, B - 1);
            const YT* src = y + (size_t)frame * n + (act ? o : 0);
            const YT v0 = src[0], v1 = src[Q], v2 = src[2 * Q], v3 = src[3 * Q];
            tb[lane * 33 + f] = q8((double)v0, sc) | (q8((double)v2, sc) << 8) | (q8((double)v1, sc) << 16) | (q8((double)v3, sc) << 24);
        }
        __syncwarp();
        const int rows = min(32, Q - g0);
        for (int r = 0; r < rows; r++) {
            const int x4 = (int)(__brev((unsigned)(g0 + r)) >> (32 - (m - 2)));
            slice[x4 * 32 + lane] = tb[r * 33 + lane];
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(32, 32) decode_sc8_kernel(const __grid_constant__ Args S) {
    const DecodeArgs& A = S.a;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x;
    const int n = A.n, m = A.m, NW = n >> 5;
    const int vs = S.vs, ntop = n >> (vs - 1);                           // bytes per frame of the levels from vs down: ntop - 8
    uint32_t* P = reinterpret_cast<uint32_t*>(smem_raw) + lane;          // level v >= vs ((n >> v) / 4 rows) at row (ntop - (n >> (v - 1))) / 4
    uint32_t* H = reinterpret_cast<uint32_t*>(smem_raw) + ((ntop - 8) >> 2) * 32 + lane;
    uint32_t* G = S.glev + (size_t)blockIdx.x * ((n - ntop) >> 2) * 32 + lane;   // levels 1 .. vs - 1: level v at row (n - (n >> (v - 1))) / 4
    auto levs = [&](int v) -> uint32_t* { return P + ((ntop - (n >> (v - 1))) >> 2) * 32; };      // v >= vs: shared memory
    auto levg = [&](int v) -> uint32_t* { return G + ((n - (n >> (v - 1))) >> 2) * 32; };         // v < vs: the block's global slice

    for (int tile = blockIdx.x; tile < S.ntiles; tile += gridDim.x) {
        const int frame = tile * 32 + lane, fr = min(frame, A.B - 1);
        const uint32_t* y0 = S.yq + (size_t)(S.fused ? blockIdx.x : tile) * (n >> 2) * 32 + lane;
        if (S.fused) {                                       // (the level region and the heap are dead between tiles: 32 x 33 words of them)
            uint32_t* tb = reinterpret_cast<uint32_t*>(smem_raw);
            uint32_t* slice = S.yq + (size_t)blockIdx.x * (n >> 2) * 32;
            if (A.y_f32) quantise_tile(reinterpret_cast<const float*>(A.y), n, m, A.B, tile, (double)S.scale, slice, tb, lane);
            else quantise_tile(reinterpret_cast<const double*>(A.y), n, m, A.B, tile, (double)S.scale, slice, tb, lane);
        }
        sct::Walk wk;
        wk.w0 = 0; wk.uacc = 0; wk.H = H; wk.bits = S.bits + (size_t)fr * NW; wk.store = frame < A.B; wk.M = m;
        bool bad = false;
#pragma unroll 1
        for (int i = 0; i < S.nops; i++) {
            const uint32_t op = __ldg(S.ops + i);
            const uint32_t type = op & 15u;
            const int v = (op >> 4) & 31, aux = (int)(op >> 13);
            const bool isg = (op >> 9) & 1u;
            if (type == sct::OP_LEVEL) {
                const int Sz = n >> v, R = Sz >> 2;                      // elements, rows
                // one inlined copy per address-space pair (global -> global, global -> shared, shared -> shared): with a
                // pointer selected at run time every access would be a generic LD / ST
                auto run = [&](const uint32_t* __restrict__ src, uint32_t* __restrict__ dst) {
                    if (isg) {
#pragma unroll 4
                        for (int r = 0; r < R; r++) {
                            const int pos = Sz + 4 * r;
                            const uint32_t nib = (Sz >= 32 ? H[(pos >> 5) * 32] >> (pos & 31) : wk.w0 >> pos) & 15u;
                            dst[r * 32] = g4(src[r * 32], src[(r + R) * 32], nib);
                        }
                    } else {
#pragma unroll 4
                        for (int r = 0; r < R; r++) dst[r * 32] = f4(src[r * 32], src[(r + R) * 32]);
                    }
                };
                if (v == 1) run(y0, levg(1));
                else if (v < vs) run(levg(v - 1), levg(v));
                else if (v == vs) run(levg(v - 1), levs(v));
                else run(levs(v - 1), levs(v));
            } else if (type == sct::OP_NODE8) {
                const uint32_t* src = levs(m - 3);
                const uint32_t p0 = src[0], p1 = src[32];
                int a[8];
#pragma unroll
                for (int j = 0; j < 4; j++) { a[j] = (int)(signed char)(p0 >> (8 * j)); a[4 + j] = (int)(signed char)(p1 >> (8 * j)); }
                const uint32_t ib = (__ldg(A.info_mask + (aux >> 2)) >> (8 * (aux & 3))) & 0xFFu;
                const uint32_t fb = (__ldg(A.frozen_val + (aux >> 2)) >> (8 * (aux & 3))) & 0xFFu;
                uint32_t ub;
                const uint32_t Tps = s32::sc_node_p<8, false, int, false>(a, ib, fb, ub, bad);
                wk.finish_block(aux, ub, Tps);
            } else if (type == sct::OP_RATE0) {
                const int nblk = 1 << v;
                if (nblk >= 4) wk.finish_zero_words(aux, nblk);
                else {
#pragma unroll 1
                    for (int b = 0; b < nblk; b++) wk.finish_block(aux + b, 0u, 0u);
                }
            }
        }
        __syncwarp();
    }
}

static size_t smem_bytes(int n, int vs) {
    const size_t walk = (size_t)(((n >> (vs - 1)) - 8) >> 2) * 128 + (size_t)(n >> 5) * 128, transpose = 32 * 33 * 4;   // (the fused quantiser's buffer)
    return walk > transpose ? walk : transpose;
}
// first shared-memory level: the smallest one that lets an SM hold 32 one-warp blocks (the hardware's limit; 6 KB per tile),
// at most m - 5 (64 values) -- n = 1024: 115 M frames/s with 10 blocks per SM (vs 2), 138 with 32 (vs 5); n = 256: vs 2; n = 2048: vs 6
static int first_smem_level(int m) {
    const char* e = getenv("PCB_8_VS");
    int vs = 2;
    while (vs < m - PCB_8_VS_DEFAULT && smem_bytes(1 << m, vs) > 6144) vs++;
    if (e) vs = atoi(e);
    if (vs < 2) vs = 2;
    if (vs > m - 3) vs = m - 3;
    return vs;
}

}  // namespace sc8

bool sc8_supported(int m) { return m >= 6 && m <= 12; }
size_t sc8_yq_bytes(int n, int B) { return (size_t)((B + 31) / 32) * 32 * (size_t)n; }
size_t sc8_glev_bytes(int n, int blocks, int vs) { return (size_t)blocks * ((n - (n >> (vs - 1))) >> 2) * 128; }

// vs: the first shared-memory level chosen for this code length (the handle keeps it: slice size and launch must agree)
cudaError_t sc8_occupancy(int m, int* bps, size_t* smem, int* vs) {
    const int n = 1 << m;
    *vs = sc8::first_smem_level(m);
    *smem = sc8::smem_bytes(n, *vs);
    if (*smem > 227 * 1024) { *bps = 0; return cudaSuccess; }
    cudaError_t e = cudaFuncSetAttribute(sc8::decode_sc8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sc8::decode_sc8_kernel, 32, *smem);
}

// the op list of a frozen pattern for this kernel (host): one op per level from the root, all-frozen subtrees pruned
int sc8_build_ops(int n, int m, const uint32_t* info_w, const uint32_t* fv_w, uint32_t* ops_out, int cap) {
    sct::Builder b;
    b.n = n; b.m = m; b.W = 1; b.info = info_w; b.fv = fv_w; b.use_rate1 = false;
    b.build_flat();
    if (ops_out) {
        if ((int)b.ops.size() > cap) return -1;
        for (size_t i = 0; i < b.ops.size(); i++) ops_out[i] = b.ops[i];
    }
    return (int)b.ops.size();
}

// a: y / y_f32 / B; qfraction: the scale is 2 << qfraction; yq: sc8_yq_bytes of scratch; bits: [B][n/32]
cudaError_t launch_decode_sc8(const DecodeArgs& a, int qfraction, const uint32_t* d_ops, int nops, uint32_t* yq, uint32_t* glev, uint32_t* bits, int blocks,
                              int sm_count, int vs, cudaStream_t st) {
    const int ntiles = (a.B + 31) / 32;
    const float scale = (float)(2 << qfraction);
    const bool fused = getenv("PCB_8_PREPASS") == nullptr;     // PCB_8_PREPASS=1: the quantiser as a pass of its own (A/B, tests)
    cudaError_t e;
    if (!fused) {
        const size_t qsm = (size_t)32 * ((a.n >> 2) + 1) * 4;
        int qpb = (int)((227 * 1024) / (qsm + 1024));             // blocks an SM holds (33 KB each at n = 1024)
        if (qpb > 8) qpb = 8;
        if (qpb < 1) qpb = 1;
        int qb = ntiles < sm_count * qpb ? ntiles : sm_count * qpb;
        if (a.y_f32) {
            e = cudaFuncSetAttribute(sc8::sc8_quantise_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qsm);
            if (e != cudaSuccess) return e;
            sc8::sc8_quantise_kernel<float><<<qb, 256, qsm, st>>>((const float*)a.y, yq, a.n, a.m, a.B, ntiles, scale);
        } else {
            e = cudaFuncSetAttribute(sc8::sc8_quantise_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qsm);
            if (e != cudaSuccess) return e;
            sc8::sc8_quantise_kernel<double><<<qb, 256, qsm, st>>>((const double*)a.y, yq, a.n, a.m, a.B, ntiles, scale);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    sc8::Args s;
    s.a = a; s.ops = d_ops; s.nops = nops; s.yq = yq; s.glev = glev; s.bits = bits; s.ntiles = ntiles; s.fused = fused ? 1 : 0; s.scale = scale;
    if (blocks > ntiles) blocks = ntiles;
    s.vs = vs;
    const size_t smem = sc8::smem_bytes(a.n, s.vs);
    e = cudaFuncSetAttribute(sc8::decode_sc8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    sc8::decode_sc8_kernel<<<blocks, 32, smem, st>>>(s);
    return cudaGetLastError();
}

}  // namespace pcb

// test hook (not part of include/polar_b200.h): the op list, for the host-side emulation in tests/
extern "C" int pcb_dbg_sc8_ops(int n, int m, const uint32_t* info_w, const uint32_t* fv_w, uint32_t* ops_out, int cap) {
    return pcb::sc8_build_ops(n, m, info_w, fv_w, ops_out, cap);
}
