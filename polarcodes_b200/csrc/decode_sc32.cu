// Min-sum SC decoder (SCDecoderMinsum.cpp:14-76), thirty-two frames per warp: lane = frame.
//
// An SC decoder is one dependent walk over the code tree; inside a frame only the f/g updates of
// one node are independent, and half of all updates belong to nodes of size <= 16.  Frames are
// independent, so a lane owns a whole frame: every update, decision and partial-sum word is
// lane-local, the warp needs no shuffle, vote or __syncwarp, and control (the frozen pattern, the
// tree walk) is warp-uniform.  The price is memory: a frame's LLR tree is n values.  Layout:
//  * the channel values are transposed once per batch into [tile of 32 frames][x][lane]
//    (sc32_transpose_kernel, HBM-bound), already gathered through the bit reversal that maps the
//    reference's adjacent pairing (SCDecoderMinsum.cpp:32-66) onto half-split pairing (DESIGN.md
//    section 3), so that every load of the decoder is one coalesced row;
//  * layers larger than 32 are never stored: every 32 phases the size-32 layer is produced through
//    fused levels (the partial sums of those levels come from the lane's bit heap); for n >= 512
//    the size-256 level is kept in a per-block global buffer that stays in L2 and is refreshed
//    every 256 phases, so the channel values are read n/256 times instead of n/32;
//  * sizes 32, 16, 8 live in shared memory as rows of 32 lanes ([row][lane]: conflict-free);
//    a size-8 subtree is decoded in registers by a fully unrolled recursion;
//  * partial sums are bit-packed per lane: sizes < 32 in one register word (size S at bits
//    [S, 2S)), larger ones in a shared-memory heap with the same rule (SCLDecoderMinsum.cpp:287-305
//    is the same combine; the SC decoder's decTree2 holds the same bits one int each).
// Arithmetic is the reference's, operation for operation, in its own precision (double; float when
// the handle was created with the fp32 precision): bit-exact decisions.
#include <cfloat>

#include "decode_common.cuh"
#include "decode_launch.h"
#include "sc_ops.cuh"

#ifndef PCB_S_FE
#define PCB_S_FE 5       // log2 of the largest stored layer: 32 (6: twice the shared memory per warp, half the fused work)
#endif
#ifndef PCB_S_BIGFUSE
#define PCB_S_BIGFUSE 1  // n > 2048: two levels per pass (the f that follows a level never re-reads it; three per pass: slower)
#endif
#ifndef PCB_S_BIGTU
#define PCB_S_BIGTU 32   // n > 2048: outputs per loop iteration of a level (twice as many loads in flight)
#endif
#ifndef PCB_S_TOPU
#define PCB_S_TOPU 8     // outputs of the fused top stage per loop iteration (16 << log2 of it channel loads in flight)
#endif

namespace pcb {
namespace s32 {

template <int M, typename T>
struct Geo {
    static constexpr int FE = PCB_S_FE, F = 1 << FE;        // largest stored layer
    static constexpr int N = 1 << M, V = M - FE, NW = N / 32;
    static constexpr int CL = 8 - FE;                        // levels between the cached size-256 level and size F
    static constexpr int PROWS = 2 * F - 8;                  // sizes F, F/2, .., 8
    __host__ __device__ static constexpr int off(int e) { return 2 * F - (2 << e); }
    static constexpr int BYTES = 32 * PROWS * (int)sizeof(T) + 32 * NW * 4;
};

struct Args {
    DecodeArgs a;
    const void* yt;          // transposed channel values
    void* scratch;           // n >= 512: 256 rows of 32 lanes per block
    uint32_t* bits;          // [B][n/32] packed decisions
    int ntiles;
    uint32_t info_w[64];     // n <= 2048
    uint32_t fv_w[64];
};

// Element x of level BASE + LEVELS from the 2^LEVELS elements of level BASE it depends on (level v
// has n >> v elements, level 0 = channel values; element x of level v pairs x and x + (n >> v) of
// level v - 1, its partial-sum bit sits at heap bit (n >> v) + x).  All loads first (one basic
// block: they are in flight together), then the levels in registers; level BASE + l is a g-level
// iff bit LEVELS - l of `path` is set (warp-uniform), chosen by select so that nothing separates
// the loads.
template <int M, int BASE, int LEVELS, int NOUT, bool EXACT, typename T>
__device__ __forceinline__ void fuse(const T* __restrict__ src, const uint32_t* H, int lane, int path, int x0, T (&out)[NOUT]) {
    constexpr int C = 1 << LEVELS, UNIT = ((1 << M) >> BASE) >> LEVELS;
    T v[NOUT][C];
#pragma unroll
    for (int o = 0; o < NOUT; o++)
#pragma unroll
        for (int t = 0; t < C; t++) v[o][t] = src[(size_t)(x0 + o + UNIT * t) * 32];
#pragma unroll
    for (int o = 0; o < NOUT; o++) {
#pragma unroll
        for (int l = 1; l <= LEVELS; l++) {
            const int half = C >> l;
            const bool isg = (path >> (LEVELS - l)) & 1;
            const int S = (1 << M) >> (BASE + l);
#pragma unroll
            for (int t = 0; t < half; t++) {
                const int xx = x0 + o + UNIT * t;
                const T a = v[o][t], b = v[o][t + half];
                const int c = isg ? (int)((H[((S + xx) >> 5) * 32 + lane] >> (xx & 31)) & 1u) : 0;
                if constexpr (EXACT) {       // three transcendentals per f: a (warp-uniform) branch is cheaper than a select
                    v[o][t] = isg ? opg<EXACT, T>(a, b, c) : opf<EXACT, T>(a, b);
                } else {
                    const T gv = opg<EXACT, T>(a, b, c), fv = opf<EXACT, T>(a, b);
                    v[o][t] = isg ? gv : fv;
                }
            }
        }
        out[o] = v[o][0];
    }
}

template <int M, typename T, bool EXACT>
__global__ void __launch_bounds__(32) decode_sc32_kernel(const __grid_constant__ Args S) {
    using G = Geo<M, T>;
    const DecodeArgs& A = S.a;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x;
    T* P = reinterpret_cast<T*>(smem_raw);
    uint32_t* H = reinterpret_cast<uint32_t*>(P + G::PROWS * 32);
    constexpr int N = G::N, TU = PCB_S_TOPU;
    T* LB = reinterpret_cast<T*>(S.scratch) + (size_t)blockIdx.x * 256 * 32 + lane;   // size-256 level of this block's tile

    for (int tile = blockIdx.x; tile < S.ntiles; tile += gridDim.x) {
        const T* yt = reinterpret_cast<const T*>(S.yt) + (size_t)tile * N * 32 + lane;
        const int frame = tile * 32 + lane;
        uint32_t* bits = S.bits + (size_t)min(frame, A.B - 1) * G::NW;   // this lane's packed decisions
        uint32_t w0 = 0, uacc = 0;
#pragma unroll 1
        for (int b64 = 0; b64 < N / G::F; ++b64) {
            // size-F layer from the channel values through V fused levels; when more than CL levels
            // are fused the size-256 level is kept in a per-block global buffer (64 KB in fp64, stays
            // in L2) and refreshed every 256 phases: n/256 passes over the channel values, not n/F
            constexpr int TUB = G::CL > 2 ? (TU >> (G::CL - 2) > 0 ? TU >> (G::CL - 2) : 1) : TU;   // same number of loads in flight
            T o[TU];
            if constexpr (G::V > G::CL) {
                if ((b64 & ((1 << G::CL) - 1)) == 0) {
#pragma unroll 1
                    for (int x = 0; x < 256; x += TU) {
                        fuse<M, 0, G::V - G::CL, TU, EXACT, T>(yt, H, lane, b64 >> G::CL, x, o);
#pragma unroll
                        for (int i = 0; i < TU; i++) LB[(size_t)(x + i) * 32] = o[i];
                    }
                }
#pragma unroll 1
                for (int j = 0; j < G::F; j += TUB) {
                    T ob[TUB];
                    fuse<M, G::V - G::CL, G::CL, TUB, EXACT, T>(LB, H, lane, b64 & ((1 << G::CL) - 1), j, ob);
#pragma unroll
                    for (int i = 0; i < TUB; i++) P[(G::off(G::FE) + j + i) * 32 + lane] = ob[i];
                }
            } else {
#pragma unroll 1
                for (int j = 0; j < G::F; j += TU) {
                    fuse<M, 0, G::V, TU, EXACT, T>(yt, H, lane, b64, j, o);
#pragma unroll
                    for (int i = 0; i < TU; i++) P[(G::off(G::FE) + j + i) * 32 + lane] = o[i];
                }
            }
#pragma unroll 1
            for (int b8 = 0; b8 < G::F / 8; ++b8) {
                const int blk = b64 * (G::F / 8) + b8;
                int e;
                if (b8 == 0) {
                    e = G::FE;                    // the f-layers below start from the size-F layer itself
                } else {
                    e = 3 + __ffs(b8) - 1;
                    const int Sz = 1 << e;
                    T* dst = P + G::off(e) * 32 + lane;
                    const T* src = P + G::off(e + 1) * 32 + lane;
                    const uint32_t cw = Sz == 32 ? H[32 + lane] : (w0 >> Sz);
#pragma unroll 4
                    for (int j = 0; j < Sz; ++j) dst[j * 32] = opg<EXACT, T>(src[j * 32], src[(j + Sz) * 32], (int)((cw >> j) & 1u));
                }
#pragma unroll 1
                for (int ee = e - 1; ee >= 3; --ee) {
                    const int Sz = 1 << ee;
                    T* dst = P + G::off(ee) * 32 + lane;
                    const T* src = P + G::off(ee + 1) * 32 + lane;
#pragma unroll 4
                    for (int j = 0; j < Sz; ++j) dst[j * 32] = opf<EXACT, T>(src[j * 32], src[(j + Sz) * 32]);
                }
                // the size-8 subtree in registers
                T a[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] = P[(G::off(3) + j) * 32 + lane];
                const uint32_t ib = (S.info_w[blk >> 2] >> (8 * (blk & 3))) & 0xFFu;
                const uint32_t fb = (S.fv_w[blk >> 2] >> (8 * (blk & 3))) & 0xFFu;
                uint32_t ub;
                uint32_t Tps = sc_node<8, EXACT, T>(a, ib, fb, ub);
                uacc |= ub << (8 * (blk & 3));
                if ((blk & 3) == 3) {
                    if (frame < A.B) bits[blk >> 2] = uacc;
                    uacc = 0;
                }
                // partial sums of the subtree that ends here, size 2^d2 (SCLDecoderMinsum.cpp:287-305)
                const int d2 = min(M, 3 + __ffs(blk + 1) - 1);
                if (d2 == 3) { w0 = (w0 & ~0xFF00u) | (Tps << 8); continue; }
                Tps = (((w0 >> 8) ^ Tps) & 0xFFu) | (Tps << 8);
                if (d2 == 4) { w0 = (w0 & 0x0000FFFFu) | (Tps << 16); continue; }
                Tps = (((w0 >> 16) ^ Tps) & 0xFFFFu) | (Tps << 16);
                if (d2 == M) continue;                     // the whole codeword: nobody reads it
                const int D = 1 << d2;
                H[(2 * D / 32 - 1) * 32 + lane] = Tps;
#pragma unroll 1
                for (int w = 32; w < D; w <<= 1) {
                    const int nw = w >> 5;
#pragma unroll 1
                    for (int kk = 0; kk < nw; kk++)
                        H[((2 * D - 2 * w) / 32 + kk) * 32 + lane] = H[(nw + kk) * 32 + lane] ^ H[((2 * D - w) / 32 + kk) * 32 + lane];
                }
            }
        }
    }
}

// ---- n > 2048 (SC min-sum at N = 32768, SURVEY config E): same lane = frame mapping, but every
// level above size 32 is stored, in a per-block global store LV[level][x][lane] (n - 32 rows of 32
// lanes), and so is the partial-sum heap (n/32 rows of 32 words).  A level is written once and read
// once or twice, always as whole coalesced rows; the decoder is DRAM-bound.  Two levels are produced
// per pass (the f-level that follows a level takes it from registers, so a level is read back only
// by the g of its right child: about 2.25 n values moved per level and frame instead of 3 n), with
// 64 loads in flight per lane.  Frozen pattern from global memory (warp-uniform loads).
struct BigArgs {
    DecodeArgs a;
    const void* yt;
    void* lv;                // per block: (n - 32) * 32 values
    uint32_t* heap;          // per block: (n / 32) * 32 words
    uint32_t* bits;
    int ntiles;
};

template <typename T, bool EXACT>
__global__ void __launch_bounds__(32) decode_sc32big_kernel(const __grid_constant__ BigArgs S) {
    const DecodeArgs& A = S.a;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x;
    const int n = A.n, m = A.m, NW = n >> 5, LT = m - 5;       // level LT has 32 elements
    T* P = reinterpret_cast<T*>(smem_raw);                       // sizes 32, 16, 8: rows 0, 32, 48
    T* LV = reinterpret_cast<T*>(S.lv) + (size_t)blockIdx.x * (n - 32) * 32 + lane;
    uint32_t* H = S.heap + (size_t)blockIdx.x * NW * 32 + lane;
    constexpr int TU = PCB_S_BIGTU;
    auto lvl = [&](int v) -> T* { return LV + (size_t)(n - (n >> (v - 1))) * 32; };   // level v >= 1, n >> v rows

    for (int tile = blockIdx.x; tile < S.ntiles; tile += gridDim.x) {
        const T* yt = reinterpret_cast<const T*>(S.yt) + (size_t)tile * n * 32 + lane;
        const int frame = tile * 32 + lane;
        uint32_t* bits = S.bits + (size_t)min(frame, A.B - 1) * NW;
        uint32_t w0 = 0, uacc = 0;
#pragma unroll 1
        for (int b32 = 0; b32 < NW; ++b32) {
            // levels v0 .. LT: the first one is a g-level (except at the very start), the rest f-levels
            const int v0 = b32 ? LT - (__ffs(b32) - 1) : 1;
#pragma unroll 1
            for (int v = v0; v <= LT;) {
                const int Sz = n >> v;
                const bool isg = b32 && v == v0;
                const T* src = v == 1 ? yt : lvl(v - 1);
#if PCB_S_BIGFUSE
                if (v < LT) {
                    // levels v and v+1 in one pass: outputs x and x + Sz/2 of level v are produced
                    // together, stored (the g of the right child reads them later) and combined on the
                    // spot into output x of level v+1, which is therefore never read back for its f
                    constexpr int TF = TU / 2;
                    const int Hs = Sz >> 1;
                    T* d1 = lvl(v);
                    T* d2 = v + 1 == LT ? P + lane : lvl(v + 1);
#pragma unroll 1
                    for (int x = 0; x < Hs; x += TF) {
                        T a0[TF], b0[TF], a1[TF], b1[TF];
#pragma unroll
                        for (int i = 0; i < TF; i++) {
                            a0[i] = src[(size_t)(x + i) * 32]; b0[i] = src[(size_t)(x + i + Sz) * 32];
                            a1[i] = src[(size_t)(x + i + Hs) * 32]; b1[i] = src[(size_t)(x + i + Hs + Sz) * 32];
                        }
                        const uint32_t cw0 = isg ? H[(size_t)((Sz + x) >> 5) * 32] >> (x & 31) : 0u;
                        const uint32_t cw1 = isg ? H[(size_t)((Sz + x + Hs) >> 5) * 32] >> (x & 31) : 0u;
#pragma unroll
                        for (int i = 0; i < TF; i++) {
                            const T o0 = isg ? opg<EXACT, T>(a0[i], b0[i], (int)((cw0 >> i) & 1u)) : opf<EXACT, T>(a0[i], b0[i]);
                            const T o1 = isg ? opg<EXACT, T>(a1[i], b1[i], (int)((cw1 >> i) & 1u)) : opf<EXACT, T>(a1[i], b1[i]);
                            d1[(size_t)(x + i) * 32] = o0;
                            d1[(size_t)(x + i + Hs) * 32] = o1;
                            d2[(size_t)(x + i) * 32] = opf<EXACT, T>(o0, o1);
                        }
                    }
                    v += 2;
                    continue;
                }
#endif
                T* dl = v == LT ? P + lane : lvl(v);          // rows of 32 lanes either way
#pragma unroll 1
                for (int x = 0; x < Sz; x += TU) {
                    T a[TU], b[TU];
#pragma unroll
                    for (int i = 0; i < TU; i++) { a[i] = src[(size_t)(x + i) * 32]; b[i] = src[(size_t)(x + i + Sz) * 32]; }
                    const uint32_t cw = isg ? H[(size_t)((Sz + x) >> 5) * 32] >> (x & 31) : 0u;
#pragma unroll
                    for (int i = 0; i < TU; i++) {
                        const T o = isg ? opg<EXACT, T>(a[i], b[i], (int)((cw >> i) & 1u)) : opf<EXACT, T>(a[i], b[i]);
                        dl[(size_t)(x + i) * 32] = o;
                    }
                }
                ++v;
            }
#pragma unroll 1
            for (int b8 = 0; b8 < 4; ++b8) {
                const int blk = b32 * 4 + b8;
                int e = 5;
                if (b8) {
                    e = 3 + __ffs(b8) - 1;
                    const int Sz = 1 << e;
                    T* dst = P + (64 - (2 << e)) * 32 + lane;
                    const T* src = P + (64 - (4 << e)) * 32 + lane;
                    const uint32_t cw = w0 >> Sz;
#pragma unroll 4
                    for (int j = 0; j < Sz; ++j) dst[j * 32] = opg<EXACT, T>(src[j * 32], src[(j + Sz) * 32], (int)((cw >> j) & 1u));
                }
#pragma unroll 1
                for (int ee = e - 1; ee >= 3; --ee) {
                    const int Sz = 1 << ee;
                    T* dst = P + (64 - (2 << ee)) * 32 + lane;
                    const T* src = P + (64 - (4 << ee)) * 32 + lane;
#pragma unroll 4
                    for (int j = 0; j < Sz; ++j) dst[j * 32] = opf<EXACT, T>(src[j * 32], src[(j + Sz) * 32]);
                }
                T a[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] = P[(48 + j) * 32 + lane];
                const uint32_t ib = (__ldg(A.info_mask + b32) >> (8 * b8)) & 0xFFu;
                const uint32_t fb = (__ldg(A.frozen_val + b32) >> (8 * b8)) & 0xFFu;
                uint32_t ub;
                uint32_t Tps = sc_node<8, EXACT, T>(a, ib, fb, ub);
                uacc |= ub << (8 * b8);
                if (b8 == 3) {
                    if (frame < A.B) bits[b32] = uacc;
                    uacc = 0;
                }
                const int d2 = min(m, 3 + __ffs(blk + 1) - 1);
                if (d2 == 3) { w0 = (w0 & ~0xFF00u) | (Tps << 8); continue; }
                Tps = (((w0 >> 8) ^ Tps) & 0xFFu) | (Tps << 8);
                if (d2 == 4) { w0 = (w0 & 0x0000FFFFu) | (Tps << 16); continue; }
                Tps = (((w0 >> 16) ^ Tps) & 0xFFFFu) | (Tps << 16);
                if (d2 == m) continue;
                const int D = 1 << d2;
                H[(size_t)(2 * D / 32 - 1) * 32] = Tps;
#pragma unroll 1
                for (int w = 32; w < D; w <<= 1) {
                    const int nw = w >> 5;
#pragma unroll 1
                    for (int kk = 0; kk < nw; kk++)
                        H[(size_t)((2 * D - 2 * w) / 32 + kk) * 32] = H[(size_t)(nw + kk) * 32] ^ H[(size_t)((2 * D - w) / 32 + kk) * 32];
                }
            }
        }
    }
}

// packed decisions -> one int per bit (the reference's layout)
__global__ void __launch_bounds__(256) sc32_expand_kernel(const uint32_t* __restrict__ bits, int* __restrict__ sol, size_t total) {
    for (size_t t = (size_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (size_t)gridDim.x * 256)
        sol[t] = (bits[t >> 5] >> (t & 31)) & 1;
}

// y[B][n] -> yt[tile][x][lane] = y[tile * 32 + lane][bitrev_m(x)], converted to T; the frames behind
// B in the last tile repeat frame B - 1.  One warp per (tile, 32 consecutive source indexes).
template <typename TI, typename T>
__global__ void __launch_bounds__(128) sc32_transpose_kernel(const TI* __restrict__ y, T* __restrict__ yt, int n, int m, int B, int ntiles,
                                                             double sigma2 /* > 0: exact SC, LLR = 2y/sigma^2 (SCDecoder.cpp:15-17, :77-79) */) {
    __shared__ T sm[4][32][33];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int chunks = n >> 5;
    const long long total = (long long)ntiles * chunks;
    for (long long wid = (long long)blockIdx.x * 4 + wib; wid < total; wid += (long long)gridDim.x * 4) {
        const int tile = (int)(wid / chunks), c = (int)(wid % chunks);
        for (int f = 0; f < 32; f++) {
            const int frame = min(tile * 32 + f, B - 1);
            T v = (T)y[(size_t)frame * n + c * 32 + lane];
            if (sigma2 > 0.) {
                v = (T)((double)v * 2. / sigma2);
                if (isinf((double)v)) v = (T)DBL_MAX;
            }
            sm[wib][f][lane] = v;
        }
        __syncwarp();
        for (int i = 0; i < 32; i++) {
            const int x = (int)(__brev((unsigned)(c * 32 + i)) >> (32 - m));
            yt[((size_t)tile * n + x) * 32 + lane] = sm[wib][lane][i];
        }
        __syncwarp();
    }
}

template <int M, typename T>
static cudaError_t launch_m(const Args& s, int blocks, bool exact, cudaStream_t st) {
    const int smem = Geo<M, T>::BYTES;
    if constexpr (sizeof(T) == 8) {
        if (exact) {
            cudaError_t e = cudaFuncSetAttribute(decode_sc32_kernel<M, T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) return e;
            decode_sc32_kernel<M, T, true><<<blocks, 32, smem, st>>>(s);
            return cudaGetLastError();
        }
    }
    cudaError_t e = cudaFuncSetAttribute(decode_sc32_kernel<M, T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    decode_sc32_kernel<M, T, false><<<blocks, 32, smem, st>>>(s);
    return cudaGetLastError();
}
template <int M, typename T>
static cudaError_t occ_m(int* bps, int* bytes) {
    const int smem = Geo<M, T>::BYTES;
    *bytes = smem;
    cudaError_t e = cudaFuncSetAttribute(decode_sc32_kernel<M, T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, decode_sc32_kernel<M, T, false>, 32, smem);
}

}  // namespace s32

// packed decisions [B][n/32] -> one int per bit [B][n]
cudaError_t launch_sc_expand(const uint32_t* bits, int* sol, size_t total, int sm_count, cudaStream_t st) {
    size_t eb = (total + 255) / 256;
    if (eb > (size_t)sm_count * 32) eb = (size_t)sm_count * 32;
    s32::sc32_expand_kernel<<<(unsigned)eb, 256, 0, st>>>(bits, sol, total);
    return cudaGetLastError();
}

bool sc32_supported(int m) { return m >= 7 && m <= 16; }
bool sc32_is_big(int m) { return m > 11; }

#define PCB_SC32_DISPATCH(FN, ...)                                                    \
    switch (m) {                                                                     \
        case 7: return fp64 ? s32::FN<7, double>(__VA_ARGS__) : s32::FN<7, float>(__VA_ARGS__);     \
        case 8: return fp64 ? s32::FN<8, double>(__VA_ARGS__) : s32::FN<8, float>(__VA_ARGS__);     \
        case 9: return fp64 ? s32::FN<9, double>(__VA_ARGS__) : s32::FN<9, float>(__VA_ARGS__);     \
        case 10: return fp64 ? s32::FN<10, double>(__VA_ARGS__) : s32::FN<10, float>(__VA_ARGS__);  \
        case 11: return fp64 ? s32::FN<11, double>(__VA_ARGS__) : s32::FN<11, float>(__VA_ARGS__);  \
    }                                                                                \
    return cudaErrorInvalidValue;

cudaError_t sc32_occupancy(int m, int fp64, int* bps, int* bytes) {
    if (sc32_is_big(m)) {
        *bytes = 56 * 32 * (fp64 ? 8 : 4);
        return fp64 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, s32::decode_sc32big_kernel<double, false>, 32, *bytes)
                    : cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, s32::decode_sc32big_kernel<float, false>, 32, *bytes);
    }
    PCB_SC32_DISPATCH(occ_m, bps, bytes)
}

// bytes of the transposed copy of a batch
size_t sc32_yt_bytes(int n, int B, int fp64) { return (size_t)((B + 31) / 32) * 32 * (size_t)n * (fp64 ? 8 : 4); }

// transposes the batch into yt (sc32_yt_bytes) and decodes it
static cudaError_t launch_dispatch(const s32::Args& s, int m, int fp64, int blocks, bool exact, cudaStream_t st) {
    PCB_SC32_DISPATCH(launch_m, s, blocks, exact, st)
}

// bytes of the per-block buffer of the size-256 level
size_t sc32_scratch_bytes(int blocks, int fp64) { return (size_t)blocks * 256 * 32 * (fp64 ? 8 : 4); }
// n > 2048: the per-block level store followed by the per-block heap
size_t sc32_big_scratch_bytes(int n, int blocks, int fp64) {
    return (size_t)blocks * ((size_t)(n - 32) * 32 * (fp64 ? 8 : 4) + (size_t)(n / 32) * 32 * 4);
}

// tmp_bits: [B][n/32] words, used when the caller wants one int per bit (a.sol) and no packed output
cudaError_t launch_decode_sc32(const DecodeArgs& a, const uint32_t* info_w, const uint32_t* fv_w, void* yt, void* scratch,
                               uint32_t* tmp_bits, int fp64, int exact, int blocks, int sm_count, cudaStream_t st) {
    s32::Args s;
    const int all_blocks = blocks;     // the scratch was sized for this many
    s.a = a;
    s.yt = yt;
    s.scratch = scratch;
    s.bits = a.sol_bits ? a.sol_bits : tmp_bits;
    s.ntiles = (a.B + 31) / 32;
    for (int i = 0; i < 64; i++) { s.info_w[i] = i < a.n / 32 ? info_w[i] : 0; s.fv_w[i] = i < a.n / 32 ? fv_w[i] : 0; }
    const double sigma2 = exact ? a.sigma * a.sigma : 0.;
    const long long warps = (long long)s.ntiles * (a.n / 32);
    int tb = (int)((warps + 3) / 4);
    if (tb > sm_count * 32) tb = sm_count * 32;
    if (fp64) {
        if (a.y_f32) s32::sc32_transpose_kernel<float, double><<<tb, 128, 0, st>>>((const float*)a.y, (double*)yt, a.n, a.m, a.B, s.ntiles, sigma2);
        else s32::sc32_transpose_kernel<double, double><<<tb, 128, 0, st>>>((const double*)a.y, (double*)yt, a.n, a.m, a.B, s.ntiles, sigma2);
    } else {
        if (a.y_f32) s32::sc32_transpose_kernel<float, float><<<tb, 128, 0, st>>>((const float*)a.y, (float*)yt, a.n, a.m, a.B, s.ntiles, sigma2);
        else s32::sc32_transpose_kernel<double, float><<<tb, 128, 0, st>>>((const double*)a.y, (float*)yt, a.n, a.m, a.B, s.ntiles, sigma2);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (blocks > s.ntiles) blocks = s.ntiles;
    if (sc32_is_big(a.m)) {
        s32::BigArgs g;
        g.a = a; g.yt = yt; g.bits = s.bits; g.ntiles = s.ntiles;
        g.lv = scratch;
        g.heap = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(scratch) + (size_t)all_blocks * (size_t)(a.n - 32) * 32 * (fp64 ? 8 : 4));
        const int smem = 56 * 32 * (fp64 ? 8 : 4);
        if (fp64) {
            if (exact) s32::decode_sc32big_kernel<double, true><<<blocks, 32, smem, st>>>(g);
            else s32::decode_sc32big_kernel<double, false><<<blocks, 32, smem, st>>>(g);
        } else {
            s32::decode_sc32big_kernel<float, false><<<blocks, 32, smem, st>>>(g);
        }
        e = cudaGetLastError();
    } else {
        e = launch_dispatch(s, a.m, fp64, blocks, exact != 0, st);
    }
    if (e != cudaSuccess) return e;
    if (a.sol) {
        const size_t total = (size_t)a.B * a.n;          // n is a multiple of 32: frame * n / 32 = frame * NW
        size_t eb = (total + 255) / 256;
        if (eb > (size_t)sm_count * 32) eb = (size_t)sm_count * 32;
        s32::sc32_expand_kernel<<<(unsigned)eb, 256, 0, st>>>(s.bits, a.sol, total);
        e = cudaGetLastError();
    }
    return e;
}

}  // namespace pcb
