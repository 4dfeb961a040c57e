// SC decoders (SCDecoderMinsum.cpp:14-76 min-sum, SCDecoder.cpp:13-86 exact), one TILE of 32 frames per
// block: lane = frame as in decode_sc32.cu, but the W warps of a block share the tile and split the rows of
// a level between them, every level a node needs is STORED (nothing is recomputed), the channel values are
// read in place (no transposed copy of the batch), and the code tree is pruned.
//
//  * Index space: element x of level v (n >> v rows) pairs rows x and x + (n >> v) of level v - 1
//    (half-split pairing on the bit-reversed channel index).  Rows are [x][lane], so
//    every access of a warp is one conflict-free row and all arithmetic of a frame stays in its lane.
//  * Levels 1 and 2 are fused: output x of level 2 needs y[4r .. 4r+3] with r = bitrev(x), 32 contiguous
//    bytes of the frame's own row (one sector per lane, read straight from the caller's batch, the next
//    group of loads in flight while the current one is consumed); level 1 is never stored.  The channel
//    values are read four times per frame (once per level-2 node).
//  * Levels VS .. m-3 (sizes .. 8) live in shared memory, levels 2 .. VS-1 in a per-block global store that
//    stays in L2 (VS is chosen so that three tiles fit an SM: while warp 0 of one tile walks its leaves the
//    other tiles use the SM); a size-8 subtree is decoded in registers; partial sums are bit-packed per lane
//    (register word + heap, the combine of SCLDecoderMinsum.cpp:287-305).
//  * Pruning (the frozen pattern is known when the code is set, so the host compiles the tree walk into a
//    list of ops the block interprets; control is uniform for the whole grid):
//      - a subtree whose positions are all frozen (values 0) needs no LLR at all: its decisions are the
//        frozen values, its partial sums are zero, and the f that would feed it is skipped;
//      - a subtree whose positions are all information bits decodes to the hard decisions of its input
//        LLRs (partial sums = signs, u = signs F^(x)s): in min-sum arithmetic every g inside adds two
//        same-signed values and every f takes a min, so no LLR of the subtree can be zero or change the
//        sign pattern PROVIDED no input is zero and no product of two inputs underflows (|alpha| >= 2^-511;
//        2^-63 in float).  A frame that violates this anywhere is put on the redo list and decoded again by
//        the reference-literal generic kernel right behind this one (DecodeArgs::redo_*).
//    What is left at (1024, 512): 62 % of the f/g updates, 43 of 128 size-8 subtrees.
//  * Ops on levels with enough rows are executed by all warps (a barrier on either side), the small ones and
//    the leaf work by warp 0 alone; the other warps read their own list (the shared ops only).
// Arithmetic is the reference's, operation for operation, in its precision: bit-exact decisions.
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "decode_common.cuh"
#include "decode_launch.h"
#include "sc_ops.cuh"
#include "sct_common.cuh"

#ifndef PCB_T_MINWARPS
#define PCB_T_MINWARPS 0   // register budget: this many warps have to fit an SM (0: the compiler's own choice)
#endif
#ifndef PCB_T_CHUNK
#define PCB_T_CHUNK 128    // bytes of a frame's row staged per step of the top op
#endif
#ifndef PCB_T_YHINT
#define PCB_T_YHINT 1      // L2 evict-first hint on the channel-value loads
#endif
#ifndef PCB_T_SERIAL_LDS
#define PCB_T_SERIAL_LDS 1 // warp 0's serial ops address the shared-memory levels directly (LDS / STS instead of generic accesses)
#endif
#ifndef PCB_T_TOPU
#define PCB_T_TOPU 4       // level-2 outputs per group of the top op (two groups of 4 * TOPU channel values in flight per lane)
#endif

namespace pcb {
namespace sct {

struct Args {
    DecodeArgs a;
    const uint32_t* ops;     // warp 0: every op
    const uint32_t* pops;    // the other warps: the ops all warps execute
    int nops, npops;
    uint32_t* bits;          // [B][n/32] packed decisions
    int ntiles;
    int vs;                  // first level in shared memory (2: all of them)
    void* glev;              // per block: levels 2 .. vs-1, then (n > 2048) the heap
};

using s32::Tiny;

// GL: some levels (and for n > 2048 the heap) live in the block's global store; all pointers are then generic
template <typename T, bool EXACT, bool GL>
struct Tile {
    T* P;                    // lane column of the shared-memory level store
    T* G;                    // GL: lane column of the global level store
    uint32_t* H;             // lane column of the heap
    int n, m, vs;
    uint32_t wk_w0;          // warp 0: copy of Walk::w0 for the g of levels with fewer than 32 rows (serial ops only)
    __device__ __forceinline__ T* lev(int v) const {         // level v: n >> v rows
        if (GL && v < vs) return G + (size_t)((n >> 1) - (n >> (v - 1))) * 32;
        const int first = GL ? vs : 2;
        return P + ((n >> (first - 1)) - (n >> (v - 1))) * 32;
    }
    // a level known to be in shared memory (v >= vs): no address-space select, the accesses compile to LDS / STS
    __device__ __forceinline__ T* lev_s(int v) const { return P + ((n >> ((GL ? vs : 2) - 1)) - (n >> (v - 1))) * 32; }
    __device__ __forceinline__ uint32_t sbit(int pos) const { return (H[(pos >> 5) * 32] << (31 - (pos & 31))) & 0x80000000u; }

    // ---- levels 1 and 2 of node q (G1: level 1 is a g-level, G2: level 2 is), optionally f to level 3 ----
    // The warp stages 128 bytes of each of its 32 frames per step with cp.async (8 lanes per frame: every global
    // access is a whole line), two steps in flight, and each lane then reads its own frame's piece: outputs
    // r .. r + RB - 1 of level 2 in bit-reversed order (rows x = bitrev(r)).
    static constexpr int STAGE_CHUNK = PCB_T_CHUNK;           // bytes of a frame's row per step (128 or 64)
    static constexpr int STAGE_ROW = STAGE_CHUNK + 16;        // + 16: LDS.128 of 32 rows is conflict-free
    static constexpr int STAGE_BYTES = 2 * 32 * STAGE_ROW;    // per warp
    template <typename YT, bool G1, bool G2>
    __device__ __forceinline__ void top2(const YT* __restrict__ ybatch, int tile, int B, bool fuse3, double sigma2, int nw, int wi, int lane,
                                         unsigned char* stage) const {
        constexpr int RB = STAGE_CHUNK / (4 * (int)sizeof(YT)); // level-2 outputs per step: 4 (double input) or 8 (float) with 128-byte chunks
        constexpr int LPF = STAGE_CHUNK / 16, FPI = 32 / LPF;  // lanes per frame, frames per cp.async instruction
        const int R = n >> 2, per = R / nw, r0 = wi * per, sh = 32 - (m - 2), nsteps = per / RB;
        T* d2 = lev(2);
        T* d3 = lev(3);
        const unsigned sbase = (unsigned)__cvta_generic_to_shared(stage);
#if PCB_T_YHINT
        // the channel values are streamed (four passes a quarter of a decode apart, far more tiles in flight than L2 holds):
        // they leave L2 first, the level store of the resident tiles stays
        uint64_t ypol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(ypol));
#endif
        auto issue = [&](int c) {
            const char* src0 = reinterpret_cast<const char*>(ybatch) + (size_t)(r0 + c * RB) * 4 * sizeof(YT) + (lane % LPF) * 16;
            const unsigned dst0 = sbase + (c & 1) * (32 * STAGE_ROW) + (lane % LPF) * 16;
#pragma unroll
            for (int j = 0; j < LPF; j++) {
                const int f = j * FPI + lane / LPF;
                const size_t row = (size_t)min(tile * 32 + f, B - 1) * n * sizeof(YT);
#if PCB_T_YHINT
                asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst0 + f * STAGE_ROW), "l"(src0 + row), "l"(ypol) : "memory");
#else
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + f * STAGE_ROW), "l"(src0 + row) : "memory");
#endif
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        issue(0);
#pragma unroll 1
        for (int c = 0; c < nsteps; c++) {
            if (c + 1 < nsteps) {
                issue(c + 1);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            __syncwarp();
            const unsigned char* mine = stage + (c & 1) * (32 * STAGE_ROW) + lane * STAGE_ROW;
            T y[RB][4];
            if constexpr (sizeof(YT) == 8) {
#pragma unroll
                for (int i = 0; i < RB; i++) {
                    const double2 a = *reinterpret_cast<const double2*>(mine + 32 * i), b = *reinterpret_cast<const double2*>(mine + 32 * i + 16);
                    y[i][0] = (T)a.x; y[i][1] = (T)a.y; y[i][2] = (T)b.x; y[i][3] = (T)b.y;
                }
            } else {
#pragma unroll
                for (int i = 0; i < RB; i++) {
                    const float4 a = *reinterpret_cast<const float4*>(mine + 16 * i);
                    y[i][0] = (T)a.x; y[i][1] = (T)a.y; y[i][2] = (T)a.z; y[i][3] = (T)a.w;
                }
            }
            if constexpr (EXACT) {                           // SCDecoder.cpp:15-17, :77-79
#pragma unroll
                for (int i = 0; i < RB; i++)
#pragma unroll
                    for (int t = 0; t < 4; t++) {
                        double v = (double)y[i][t] * 2. / sigma2;
                        if (isinf(v)) v = DBL_MAX;
                        y[i][t] = (T)v;
                    }
            }
            const int r = r0 + c * RB;
            T o[RB];
#pragma unroll
            for (int i = 0; i < RB; i++) {
                const int x = (int)(__brev((unsigned)(r + i)) >> sh);
                T la, lb;
                if constexpr (G1) {
                    la = s32::opg2<EXACT, T>(y[i][0], y[i][1], sbit((n >> 1) + x));
                    lb = s32::opg2<EXACT, T>(y[i][2], y[i][3], sbit((n >> 1) + x + (n >> 2)));
                } else {
                    la = s32::opf2<EXACT, T>(y[i][0], y[i][1]);
                    lb = s32::opf2<EXACT, T>(y[i][2], y[i][3]);
                }
                if constexpr (G2) o[i] = s32::opg2<EXACT, T>(la, lb, sbit((n >> 2) + x));
                else o[i] = s32::opf2<EXACT, T>(la, lb);
                d2[(size_t)x * 32] = o[i];
            }
            if (fuse3) {                                     // rows x and x + n/8 of level 2 sit at r and r + 1
#pragma unroll
                for (int i = 0; i < RB; i += 2) d3[(size_t)(__brev((unsigned)(r + i)) >> sh) * 32] = s32::opf2<EXACT, T>(o[i], o[i + 1]);
            }
            __syncwarp();                                    // this buffer takes step c + 2 next
        }
    }
    template <typename YT>
    __device__ __forceinline__ void top2q(const YT* __restrict__ ybatch, int tile, int B, int q, bool fuse3, double sigma2, int nw, int wi, int lane,
                                          unsigned char* stage) const {
        if (q == 0) top2<YT, false, false>(ybatch, tile, B, fuse3, sigma2, nw, wi, lane, stage);
        else if (q == 1) top2<YT, false, true>(ybatch, tile, B, fuse3, sigma2, nw, wi, lane, stage);
        else if (q == 2) top2<YT, true, false>(ybatch, tile, B, fuse3, sigma2, nw, wi, lane, stage);
        else top2<YT, true, true>(ybatch, tile, B, fuse3, sigma2, nw, wi, lane, stage);
    }

    // ---- level v (S = n >> v rows) from level v - 1; fuse: level v + 1 = f(level v) in the same pass ----
    // SM: source and destinations are shared-memory levels (v - 1 >= vs)
    template <bool ISG, bool SM>
    __device__ __forceinline__ void level_t(int v, bool fuse, int nw, int wi) const {
        constexpr int U = PCB_T_U;
        const int S = n >> v;
        const T* src = SM ? lev_s(v - 1) : lev(v - 1);
        T* dst = SM ? lev_s(v) : lev(v);
        if (fuse) {
            T* dst2 = SM ? lev_s(v + 1) : lev(v + 1);
            const int Hs = S >> 1, per = Hs / nw, x0 = wi * per;      // (the host makes per a multiple of U)
            const T* s0 = src + (size_t)x0 * 32;
            T* q0 = dst + (size_t)x0 * 32;
            T* q2 = dst2 + (size_t)x0 * 32;
            const size_t oS = (size_t)S * 32, oH = (size_t)Hs * 32;
#pragma unroll 1
            for (int x = x0; x < x0 + per; x += U, s0 += U * 32, q0 += U * 32, q2 += U * 32) {
                // (issuing the loads of the next iteration before the arithmetic of this one doubles the registers: 50 instead of 77 M frames/s)
                T a0[U], b0[U], a1[U], b1[U];
#pragma unroll
                for (int i = 0; i < U; i++) {
                    a0[i] = s0[i * 32]; b0[i] = s0[oS + i * 32];
                    a1[i] = s0[oH + i * 32]; b1[i] = s0[oH + oS + i * 32];
                }
                uint32_t c0 = 0, c1 = 0;                             // the bit of row x + i at bit 32 - U + i
                if constexpr (ISG) {
                    if (S >= 32) { c0 = H[((S + x) >> 5) * 32] << (32 - U - (x & 31)); c1 = H[((S + x + Hs) >> 5) * 32] << (32 - U - ((x + Hs) & 31)); }
                    else { c0 = wk_w0 << (32 - U - (S + x)); c1 = wk_w0 << (32 - U - (S + x + Hs)); }
                }
#pragma unroll
                for (int i = 0; i < U; i++) {
                    T o0, o1;
                    if constexpr (ISG) {
                        o0 = s32::opg2<EXACT, T>(a0[i], b0[i], (c0 << (U - 1 - i)) & 0x80000000u);
                        o1 = s32::opg2<EXACT, T>(a1[i], b1[i], (c1 << (U - 1 - i)) & 0x80000000u);
                    } else {
                        o0 = s32::opf2<EXACT, T>(a0[i], b0[i]);
                        o1 = s32::opf2<EXACT, T>(a1[i], b1[i]);
                    }
                    q0[i * 32] = o0;
                    q0[oH + i * 32] = o1;
                    q2[i * 32] = s32::opf2<EXACT, T>(o0, o1);
                }
            }
            return;
        }
        const int per = S / nw, x0 = wi * per;
        const T* s0 = src + (size_t)x0 * 32;
        T* q0 = dst + (size_t)x0 * 32;
        const size_t oS = (size_t)S * 32;
#pragma unroll 1
        for (int x = x0; x < x0 + per; x += U, s0 += U * 32, q0 += U * 32) {
            T a[U], b[U];
#pragma unroll
            for (int i = 0; i < U; i++) { a[i] = s0[i * 32]; b[i] = s0[oS + i * 32]; }
            uint32_t c = 0;
            if constexpr (ISG) c = S >= 32 ? H[((S + x) >> 5) * 32] << (32 - U - (x & 31)) : wk_w0 << (32 - U - (S + x));
#pragma unroll
            for (int i = 0; i < U; i++) {
                if constexpr (ISG) q0[i * 32] = s32::opg2<EXACT, T>(a[i], b[i], (c << (U - 1 - i)) & 0x80000000u);
                else q0[i * 32] = s32::opf2<EXACT, T>(a[i], b[i]);
            }
        }
    }
    __device__ __forceinline__ void level(int v, bool isg, bool fuse, int nw, int wi) const {
        if (GL && PCB_T_SERIAL_LDS && v - 1 >= vs) {             // shared memory to shared memory
            if (isg) level_t<true, true>(v, fuse, nw, wi);
            else level_t<false, true>(v, fuse, nw, wi);
        } else {
            if (isg) level_t<true, false>(v, fuse, nw, wi);
            else level_t<false, false>(v, fuse, nw, wi);
        }
    }
};

// HG: the heap is in the global store too (n > 2048); otherwise its accesses compile to LDS / STS.  The n > 2048 kernels are
// bound by DRAM latency and want 20 warps per SM (at most 96 registers; 24 for fp32); the others run faster with the registers
// the compiler takes (122 for fp64, four blocks of four warps per SM) than with a fifth block.
template <typename T, bool EXACT, int W, bool GL, bool HG>
__global__ void __launch_bounds__(32 * W, EXACT ? 0 : (HG ? (sizeof(T) == 4 ? 24 : 20) : PCB_T_MINWARPS) / W) decode_sct_kernel(const __grid_constant__ Args S) {
    const DecodeArgs& A = S.a;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = A.n, m = A.m, NW = n >> 5;
    Tile<T, EXACT, GL> t;
    t.n = n; t.m = m; t.vs = GL ? S.vs : 2;
    t.wk_w0 = 0;
    const int srows = (n >> (t.vs - 1)) - 8;                   // rows of the shared-memory levels
    t.P = reinterpret_cast<T*>(smem_raw) + lane;
    t.G = nullptr;
    // [levels vs .. m-3 | staging area of the top op, one slice per warp][heap (n <= 2048)][staging area when it cannot alias the levels]
    // With vs >= 4 the top op writes no shared-memory level (levels 2, 3 are global) and every shared-memory level is dead
    // while it runs (recomputed before its next use), so the staging slices alias the level region.
    const bool alias = GL && t.vs >= 4;
    size_t region = (size_t)srows * 32 * sizeof(T);
    if (alias && region < (size_t)W * Tile<T, EXACT, GL>::STAGE_BYTES) region = (size_t)W * Tile<T, EXACT, GL>::STAGE_BYTES;
    const size_t heap_bytes = HG ? 0 : (size_t)NW * 32 * 4;
    t.H = reinterpret_cast<uint32_t*>(smem_raw + region) + lane;
    unsigned char* stage = smem_raw + (alias ? 0 : region + heap_bytes) + (size_t)warp * Tile<T, EXACT, GL>::STAGE_BYTES;
    if constexpr (GL) {
        const size_t grows = (size_t)(n >> 1) - (n >> (t.vs - 1));
        const size_t per_block = grows * 32 * sizeof(T) + (HG ? (size_t)NW * 32 * 4 : 0);
        unsigned char* gb = reinterpret_cast<unsigned char*>(S.glev) + (size_t)blockIdx.x * per_block;
        t.G = reinterpret_cast<T*>(gb) + lane;
        if constexpr (HG) t.H = reinterpret_cast<uint32_t*>(gb + grows * 32 * sizeof(T)) + lane;
    }
    const double sigma2 = A.sigma * A.sigma;
    const uint32_t* ops = warp == 0 ? S.ops : S.pops;
    const int nops = warp == 0 ? S.nops : S.npops;

    for (int tile = blockIdx.x; tile < S.ntiles; tile += gridDim.x) {
        const int frame = tile * 32 + lane, fr = min(frame, A.B - 1);
        Walk wk;
        wk.w0 = 0; wk.uacc = 0; wk.H = t.H; wk.bits = S.bits + (size_t)fr * NW; wk.store = frame < A.B; wk.M = m;
        bool bad = false;
#pragma unroll 1
        for (int i = 0; i < nops; i++) {
            const uint32_t op = __ldg(ops + i);
            const uint32_t type = op & 15u;
            const int v = (op >> 4) & 31, aux = (int)(op >> 13);
            const bool isg = (op >> 9) & 1u, fuse = (op >> 10) & 1u, par = (op >> 11) & 1u, syncb = (op >> 12) & 1u;
            if (type == OP_TOP2) {
                if (syncb) __syncthreads();
                if (A.y_f32) t.template top2q<float>(reinterpret_cast<const float*>(A.y), tile, A.B, aux, fuse, sigma2, W, warp, lane, stage);
                else t.template top2q<double>(reinterpret_cast<const double*>(A.y), tile, A.B, aux, fuse, sigma2, W, warp, lane, stage);
                __syncthreads();
            } else if (type == OP_LEVEL) {
                if (par) {
                    if (syncb) __syncthreads();
                    t.level(v, isg, fuse, W, warp);
                    __syncthreads();
                } else {                                   // (warp 0 only: the other warps' list has no such op)
                    t.wk_w0 = wk.w0;
                    t.level(v, isg, fuse, 1, 0);
                }
            } else if (type == OP_NODE8) {
                T a[8];
                const T* src = PCB_T_SERIAL_LDS ? t.lev_s(m - 3) : t.lev(m - 3);
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] = src[j * 32];
                const uint32_t ib = (__ldg(A.info_mask + (aux >> 2)) >> (8 * (aux & 3))) & 0xFFu;
                const uint32_t fb = (__ldg(A.frozen_val + (aux >> 2)) >> (8 * (aux & 3))) & 0xFFu;
                uint32_t ub;
                const uint32_t Tps = s32::sc_node_p<8, EXACT, T>(a, ib, fb, ub, bad);
                wk.finish_block(aux, ub, Tps);
            } else if (type == OP_RATE0) {                 // v = log2 of the number of 8-phase blocks
                const int nblk = 1 << v;
                if (nblk >= 4) wk.finish_zero_words(aux, nblk);
                else {
#pragma unroll 1
                    for (int b = 0; b < nblk; b++) wk.finish_block(aux + b, 0u, 0u);
                }
            } else {                                       // OP_RATE1: v = log2 of the node size (8, 16 or 32)
                const int Sz = 1 << v;
                uint32_t beta = 0;
                bool ok = true;
                auto hard = [&](const T* src) {
#pragma unroll 4
                    for (int j = 0; j < Sz; j++) {
                        const T al = src[(size_t)j * 32];
                        ok = ok && Tiny<T>::ok(al);
                        beta |= (al > (T)0 ? 0u : 1u) << j;
                    }
                };
                if (GL && PCB_T_SERIAL_LDS && m - v >= t.vs) hard(t.lev_s(m - v));
                else hard(t.lev(m - v));
                bad = bad || !ok;
                uint32_t u = beta;                         // u = beta F^(x)v, packed
                u ^= (u >> 1) & 0x55555555u; u ^= (u >> 2) & 0x33333333u; u ^= (u >> 4) & 0x0F0F0F0Fu;
                if (Sz >= 16) u ^= (u >> 8) & 0x00FF00FFu;
                if (Sz >= 32) u ^= (u >> 16) & 0x0000FFFFu;
#pragma unroll 1
                for (int b = 0; b < Sz / 8; b++) {
                    const uint32_t ub = (u >> (8 * b)) & 0xFFu;
                    wk.finish_block(aux + b, ub, xform8(ub));
                }
            }
        }
        if (warp == 0 && bad && wk.store && A.redo_count) A.redo_list[atomicAdd(A.redo_count, 1)] = frame;
        __syncthreads();                                   // the next tile reuses the level store and the heap
    }
}

template <typename T>
static size_t smem_bytes(int n, int vs, int W) {      // (the layout of decode_sct_kernel)
    const size_t levels = (size_t)((n >> (vs - 1)) - 8) * 32 * sizeof(T), heap = n > 2048 ? 0 : (size_t)(n >> 5) * 32 * 4;
    const size_t stage = (size_t)W * (2 * 32 * (PCB_T_CHUNK + 16));
    if (vs >= 4) return (levels > stage ? levels : stage) + heap;
    return levels + heap + stage;
}
template <typename T>
static size_t glob_bytes(int n, int vs) {
    return (size_t)((n >> 1) - (n >> (vs - 1))) * 32 * sizeof(T) + (n > 2048 ? (size_t)(n >> 5) * 32 * 4 : 0);
}

template <typename T, bool EXACT, int W, bool GL, bool HG>
static cudaError_t attr(int n, int vs, int* bps) {
    const size_t smem = smem_bytes<T>(n, vs, W);
    cudaError_t e = cudaFuncSetAttribute(decode_sct_kernel<T, EXACT, W, GL, HG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, decode_sct_kernel<T, EXACT, W, GL, HG>, 32 * W, smem);
}
template <typename T, bool EXACT, int W, bool GL, bool HG>
static cudaError_t go(const Args& s, int blocks, cudaStream_t st) {
    const size_t smem = smem_bytes<T>(s.a.n, s.vs, W);
    cudaError_t e = cudaFuncSetAttribute(decode_sct_kernel<T, EXACT, W, GL, HG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   // (the occupancy probe tries several sizes)
    if (e != cudaSuccess) return e;
    decode_sct_kernel<T, EXACT, W, GL, HG><<<blocks, 32 * W, smem, st>>>(s);
    return cudaGetLastError();
}

}  // namespace sct

// W: 1, 2, 4 or 8 warps per tile.  vs: first level in shared memory (2 = all).
#define PCB_SCT_G(FN, WW, GG, HH, ...)                                                                                   \
    {                                                                                                                     \
        if (exact) return sct::FN<double, true, WW, GG, HH>(__VA_ARGS__);                                                 \
        return fp64 ? sct::FN<double, false, WW, GG, HH>(__VA_ARGS__) : sct::FN<float, false, WW, GG, HH>(__VA_ARGS__);   \
    }
#define PCB_SCT_W(FN, WW, ...)                                                                                            \
    {                                                                                                                     \
        if (vs > 2 && n > 2048) PCB_SCT_G(FN, WW, true, true, __VA_ARGS__)                                                \
        if (vs > 2) PCB_SCT_G(FN, WW, true, false, __VA_ARGS__)                                                           \
        PCB_SCT_G(FN, WW, false, false, __VA_ARGS__)                                                                      \
    }
#define PCB_SCT_DISPATCH(FN, ...)                   \
    if (W == 8) PCB_SCT_W(FN, 8, __VA_ARGS__)       \
    if (W == 2) PCB_SCT_W(FN, 2, __VA_ARGS__)       \
    if (W == 1) PCB_SCT_W(FN, 1, __VA_ARGS__)       \
    PCB_SCT_W(FN, 4, __VA_ARGS__)

bool sct_supported(int m) { return m >= 7 && m <= 16; }

// shared memory per block, global store per block, resident blocks per SM for (n, vs, W)
cudaError_t sct_occupancy(int m, int fp64, int exact, int W, int vs, int* bps, size_t* smem, size_t* glob) {
    const int n = 1 << m;
    *smem = fp64 ? sct::smem_bytes<double>(n, vs, W) : sct::smem_bytes<float>(n, vs, W);
    *glob = fp64 ? sct::glob_bytes<double>(n, vs) : sct::glob_bytes<float>(n, vs);
    if (*smem > 227 * 1024) { *bps = 0; return cudaSuccess; }
    PCB_SCT_DISPATCH(attr, n, vs, bps)
}

// the two op lists of a frozen pattern (host): every op (warp 0), the ops all warps execute; null pointers to size them
int sct_build_ops(int n, int m, int W, const uint32_t* info_w, const uint32_t* fv_w, uint32_t* ops_out, int cap, uint32_t* pops_out, int pcap,
                  int* npops) {
    sct::Builder b;
    b.n = n; b.m = m; b.W = W; b.info = info_w; b.fv = fv_w;
    b.build();
    if (npops) *npops = (int)b.pops.size();
    if (ops_out) {
        if ((int)b.ops.size() > cap) return -1;
        for (size_t i = 0; i < b.ops.size(); i++) ops_out[i] = b.ops[i];
    }
    if (pops_out) {
        if ((int)b.pops.size() > pcap) return -1;
        for (size_t i = 0; i < b.pops.size(); i++) pops_out[i] = b.pops[i];
    }
    return (int)b.ops.size();
}

// a: y / y_f32 / B / redo_count + redo_list (device; the count is zeroed by the caller) as for the list kernel;
// d_ops / d_pops: device copies of the op lists; bits: [B][n/32] packed decisions; glev: blocks * glob bytes
cudaError_t launch_decode_sct(const DecodeArgs& a, const uint32_t* d_ops, int nops, const uint32_t* d_pops, int npops, uint32_t* bits, void* glev,
                              int fp64, int exact, int W, int vs, int blocks, cudaStream_t st) {
    sct::Args s;
    s.a = a; s.ops = d_ops; s.nops = nops; s.pops = d_pops; s.npops = npops; s.bits = bits; s.ntiles = (a.B + 31) / 32;
    s.vs = vs; s.glev = glev;
    if (blocks > s.ntiles) blocks = s.ntiles;
    const int n = a.n;
    static const bool dbg = getenv("PCB_DEBUG_SCT") != nullptr;
    if (dbg) fprintf(stderr, "sct launch: n %d W %d vs %d blocks %d nops %d npops %d fp64 %d exact %d pending error: %s\n", a.n, W, vs, blocks, nops, npops, fp64, exact,
                     cudaGetErrorString(cudaPeekAtLastError()));
    PCB_SCT_DISPATCH(go, s, blocks, st)
}

}  // namespace pcb

// test hook (not part of include/polar_b200.h): the op list of a frozen pattern, for the host-side emulation in tests/
extern "C" int pcb_dbg_sct_ops(int n, int m, const uint32_t* info_w, const uint32_t* fv_w, uint32_t* ops_out, int cap) {
    return pcb::sct_build_ops(n, m, 4, info_w, fv_w, ops_out, cap, nullptr, 0, nullptr);
}
