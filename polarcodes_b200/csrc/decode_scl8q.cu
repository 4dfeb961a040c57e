// SCL min-sum decoder with ONE LANE PER PATH (SCLDecoderMinsum.cpp:225-478 semantics, bit-exact
// with decode_generic.cu), templated on the list size L = 8, 16, 32: a frame takes L lanes and a
// warp decodes 32 / L frames in lockstep, lane = f*L + p with f the frame of the group and p the
// path slot.  One lane owns one path of one frame from the channel values to the CRC.  The
// comments below are written for the headline case L = 8 (four frames per warp, a "quad").
//
// Why this mapping: every frame of a batch shares the frozen pattern, so the
// control flow of the n phases is identical for all frames and a warp can run several of them
// in lockstep.  All per-path scalar work (leaf LLRs of the size-4/2/1 layers, metric update,
// fork bookkeeping, partial-sum words, CRC) then uses 32 of 32 lanes instead of L of 32, and
// every address computation / loop / branch of the upper layers is paid once per warp.
//
//  * LLR layers LS..m-3 (sizes .. 8) live in shared memory as [frame][layer][slot][j] with a
//    slot stride of size+4 floats: the 8 lanes of a quarter-warp LDS.128 hit 8 distinct 16-byte
//    bank groups for any permutation of 8 source slots (lazy reads).  For n >= 1024 the layers
//    3, 4 (and 5 for n = 2048 with L >= 16) live in a per-block global scratch that stays in L2.
//  * layers 0, 1, 2 are never stored: layer 3 is computed from the channel values with three
//    fused levels every n/8 phases.  In the first half of the frame layer 1 = f(y, y') does not
//    depend on the path, so the lanes of a frame compute it once, cooperatively, while staging
//    (the staging area is the region of the shared-memory layers, dead at those phases); in the
//    second half the channel values are staged with cp.async two chunks ahead of the arithmetic.
//  * layers of size 4, 2, 1 are registers of the owning lane; a clone copies the live ones with
//    shuffles inside its group of L lanes.
//  * lazy copy as in decode_scl8.cu: slot p writes its own arrays, a per-path word holds the slot
//    to read each layer from (log2 L bits per layer).
//  * rate-0, repetition and rate-1 nodes of size 4 (rate-1 also of size 8) have routines that
//    produce the results of the reference's phase-by-phase loop without running it.
#include "decode_common.cuh"
#include "decode_launch.h"
#define PCB_PHILOX_COMPACT 1
#include "gen_common.cuh"

// Compile-time switches, each measured on the same B200 box (tools/variant.sh + tools/ab.sh):
// register chains cover the last PCB_Q_CHAIN + 1 stored layers (sizes 8 .. 8 << PCB_Q_CHAIN)
#ifndef PCB_Q_CHAIN
#define PCB_Q_CHAIN 1
#endif
// PCB_Q_GL4: layer 4 lives in the global scratch next to layer 3, which leaves 13 KB of shared
// memory per quad and 16 resident quads per SM instead of 10 (25.5 vs 22.0 Mframes/s).  The
// scratch is then 24 KB per quad, 57 MB for the whole grid: it has to stay in L2, so ...
#ifndef PCB_Q_L5N
#define PCB_Q_L5N 8   // float4 column pairs per iteration of the layer-5 loop (GL4): 16 loads in flight (4: -0.8 %, 2: -2.6 %)
#endif
#ifndef PCB_Q_L4N
#define PCB_Q_L4N 2   // column pairs per iteration of the layer-4 g loop, 4 float4 loads each (4 pairs: 117 registers, -0.8 %)
#endif
#ifndef PCB_Q_STU
#define PCB_Q_STU 4    // unroll factors of the two staging loops of the top stage
#endif
#ifndef PCB_Q_STG
#define PCB_Q_STG 16
#endif
#ifndef PCB_Q_FADD2
#define PCB_Q_FADD2 1   // g on a float4 with two packed adds
#endif
#ifndef PCB_Q_TOPU
#define PCB_Q_TOPU 1   // unroll factor of the top-stage column loop
#endif
#ifndef PCB_Q_GL4
#define PCB_Q_GL4 1
#endif
// g-layer 1 of the top stage staged with cp.async, two chunks ahead of the arithmetic
#ifndef PCB_Q_GL5
#define PCB_Q_GL5 1
#endif
#ifndef PCB_Q_GH
#define PCB_Q_GH 1         // N = 2048, 16/32 paths: the two largest partial-sum layers in the global scratch
#endif
#ifndef PCB_Q_GH_SMALL
#define PCB_Q_GH_SMALL 0   // the same for the N = 1024 variants
#endif
#ifndef PCB_Q_GH_M10S
#define PCB_Q_GH_M10S 0    // N = 1024, 2 or 4 paths: measured L = 4 +2 % (12 warps instead of 10), L = 2 -5 %: off
#endif
#ifndef PCB_Q_GH_M11
#define PCB_Q_GH_M11 1     // N = 2048, 2..8 paths (the shipped conf.txt): 20 instead of 26 KB of shared memory per warp, 11 resident warps instead of 8
#endif
#ifndef PCB_Q_RL3
#define PCB_Q_RL3 1        // N = 2048, 16/32 paths: layer 3 is never stored -- the g of layer 4 recomputes it from the channel values
#endif
#ifndef PCB_Q_RL3_SMALL
#define PCB_Q_RL3_SMALL 0  // the same for the other variants with layer 4 in the scratch (L = 8, N = 1024: the headline)
#endif
#ifndef PCB_Q_R1FAST
#define PCB_Q_R1FAST 1     // rate-1 nodes of size 4 and 8 from the signs and the smallest magnitude of their inputs
#endif
#ifndef PCB_Q_MERGE
#define PCB_Q_MERGE 1
#endif
#ifndef PCB_Q_MERGE_SMALL
#define PCB_Q_MERGE_SMALL 0
#endif
// PCB_Q_MW: all resident warps of an SM in ONE block (each warp still decodes its own frames and
// never synchronises with the others): the 1 KB of shared memory the system reserves per block is
// paid once, which makes room for a 17th warp at n = 1024, L = 8 -- where it loses 0.9 % (the
// compiler keeps 88 instead of 106 registers under the 544-thread bound); with one frame per warp
// (L = 32, 18 warps either way) it gains 3 %, so it is on for L >= 16 only.
#ifndef PCB_Q_MW
#define PCB_Q_MW 0
#endif
#ifndef PCB_Q_MW_BIG
#define PCB_Q_MW_BIG 1
#endif
#ifndef PCB_Q_WCAP
#define PCB_Q_WCAP 20      // most warps of such a block (20 warps: 102 registers each, all of the shared memory; 18: -1 %)
#endif
#ifndef PCB_Q_CPA
#define PCB_Q_CPA 1
#endif
#ifndef PCB_Q_PROU
#define PCB_Q_PROU 1   // unroll factor of the prologue loops (narrowing / range check): code size beats loads in flight (1: 35.9, 2: 35.8, 4: 35.4 M frames/s)
#endif
#ifndef PCB_Q_PF64
#define PCB_Q_PF64 0   // L2 prefetch of the next quad's double rows, late in the current quad: 33.5 vs 35.7 M frames/s, off
#endif
#ifndef PCB_Q_CPU
#define PCB_Q_CPU 4    // unroll factor of the cp.async issue loop
#endif
// ... (caching layer 2 of an even top-stage event in the scratch for the odd event that follows,
// +32 KB per quad, was a 4 % gain with 10 quads per SM and a 15 % loss with 16: removed.)

namespace pcb {

namespace q8 {

constexpr unsigned FULL = 0xffffffffu;

template <int M, int LL>
struct Geo {
    static_assert(LL == 2 || LL == 4 || LL == 8 || LL == 16 || LL == 32, "list size");
    static constexpr int FPW = 32 / LL;     // frames per warp: lane = frame * LL + path slot
    static constexpr int LB = LL == 2 ? 1 : LL == 4 ? 2 : LL == 8 ? 3 : LL == 16 ? 4 : 5;   // bits of a slot number
    static constexpr unsigned GM = LL == 32 ? 0xffffffffu : (1u << LL) - 1;   // lane mask of one frame (shifted down)
    static constexpr int N = 1 << M;
    static constexpr int LTOP = 3;          // first stored LLR layer (size N/8)
    static constexpr int LBOT = M - 3;      // last stored LLR layer (size 8); sizes 4,2,1 are registers
    // N >= 1024: layer 3 (half of all stored LLRs) lives in a per-block global scratch that stays in
    // L2, which doubles the resident frames per SM; it is written and read 8 times per frame.
    static constexpr bool GL3 = M >= 10;
    // optional experiment: layer 4 there too (see PCB_Q_GL4)
    static constexpr bool GL4 = GL3 && PCB_Q_GL4;
    // N = 2048 with 16 or 32 paths per frame: layer 5 (size 64) as well, 17 KB instead of 26 KB of
    // shared memory per warp (12 resident warps per SM instead of 8; the throughput follows the
    // number of resident warps).  With 8 paths the staging area of the top stage would not fit.
    static constexpr bool GL5 = GL4 && M >= 11 && LL >= 16 && PCB_Q_GL5;
    // RL3 (with GL5): layer 3 (half of the scratch, 32 KB per frame at L = 32) is not stored at all.  The top stage
    // produces layers 3 and 4 together from the channel values; at the g of layer 4 it runs again (same partial sums of
    // layers 1-3: they only change at top-stage boundaries, and a clone inherits the slot fields that find them) and
    // combines the two halves with g instead of f.  More arithmetic, far less scratch: what is left fits L2.
    static constexpr bool RL3 = GL4 && (GL5 ? PCB_Q_RL3 : PCB_Q_RL3_SMALL);
    static constexpr int LS = GL5 ? 6 : GL4 ? 5 : GL3 ? 4 : 3;  // first LLR layer in shared memory
    static constexpr int NW = N / 32;       // words of a full-length bit array
    // Partial-sum heap of a slot: the layer of size S at words [S/32, 2S/32).  GH: the two largest
    // layers (words >= SPLITW, three quarters of the heap) live in the global scratch as
    // [word][slot]; they are touched at the top-stage events and by the rare big combines only.
    static constexpr bool GH = M >= 10 && (M >= 11 ? (LL >= 16 ? PCB_Q_GH : PCB_Q_GH_M11) : LL <= 4 ? PCB_Q_GH_M10S : PCB_Q_GH_SMALL);
    static constexpr int SPLITW = GH ? NW / 4 : (1 << 20);
    static constexpr int HS = GH ? SPLITW + 1 : NW + 1;       // slot stride of the shared-memory part
    static constexpr int XHS = NW + 1;      // slot stride of the final codeword estimates
    static constexpr int Q3 = N / 32;       // float4s of one layer-3 array
    __host__ __device__ static constexpr int size(int lam) { return N >> lam; }
    __host__ __device__ static constexpr int stride(int lam) { return size(lam) + 4; }
    __host__ __device__ static constexpr int off(int lam) {
        int o = 0;
        for (int l = LS; l < lam; ++l) o += LL * stride(l);
        return o;
    }
    static constexpr int LFLOATS = off(LBOT + 1);             // the shared-memory LLR layers
    // staging area of the top stage = the layers below the one it writes (dead at those phases): one frame's
    // layer 1 (N/2 floats) or, for N < 1024, its channel values.  With 2 or 4 paths per frame the layers are
    // smaller than that and the region is extended (a frame then costs about what it costs with 8 paths).
    static constexpr int STAGE_OFF = GL4 ? 0 : off(LS + 1);
    // CHUNKF (2 or 4 paths, N >= 1024): layer 1 is never staged whole; the channel values go through the
    // chunked cp.async staging at every top-stage event (512 floats per frame) and an f-layer 1 is
    // computed from them per path.  That costs some redundant f's and saves the shared memory that
    // decides how many warps are resident.
    // Only where it buys resident warps: N = 2048 (8 instead of 5 per SM at L = 4: 17.5 vs 14.8 M frames/s); at
    // N = 1024 the frame needs 512 floats either way and the redundant f's cost 4 % (L = 4) to 15 % (L = 2).
    static constexpr bool CHUNKF = GL3 && M >= 11 && LL <= 4 && PCB_Q_CPA;
    static constexpr int STAGE_NEED = CHUNKF ? 512 : GL3 ? N / 2 : N;
    static constexpr int PFLOATS = LFLOATS - STAGE_OFF >= STAGE_NEED ? LFLOATS : STAGE_OFF + STAGE_NEED;
    static_assert(LL <= 4 || PFLOATS == LFLOATS, "staging area too small");
    static constexpr int XW = LL * XHS;                      // final codeword estimates (alias the LLR region)
    static_assert(XW <= PFLOATS, "final codeword estimates do not fit the LLR region");
    // select by merging two sorted lists of L (fork_slow): 4L words of key scratch instead of 2L
    static constexpr bool MERGE = LL >= 16 ? PCB_Q_MERGE : PCB_Q_MERGE_SMALL;
    static constexpr int OFF_H = PFLOATS;
    static constexpr int OFF_KEYS = (PFLOATS + LL * HS + 3) & ~3;
    static constexpr int FRAME_WORDS = (OFF_KEYS + (MERGE ? 4 : 2) * LL + 3) & ~3;   // LLRs, heaps, candidate keys
    static constexpr int TOTAL = FPW * FRAME_WORDS * 4;
    static constexpr int L3_BLOCK_FLOAT4 = GL3 && !RL3 ? FPW * Q3 * LL : 0;   // global scratch per block: [frame][k][slot] float4
    static constexpr int L4_BLOCK_FLOAT4 = GL4 ? FPW * (Q3 / 2) * LL : 0;
    static constexpr int L5_BLOCK_FLOAT4 = GL5 ? FPW * (Q3 / 4) * LL : 0;
    static constexpr int HG_BLOCK_FLOAT4 = GH ? FPW * (NW - NW / 4) * LL / 4 : 0;
    static constexpr int SCRATCH_BLOCK_FLOAT4 = L3_BLOCK_FLOAT4 + L4_BLOCK_FLOAT4 + L5_BLOCK_FLOAT4 + HG_BLOCK_FLOAT4;
    // one slot number per layer: LLR layers LTOP..LBOT in pidx, partial-sum layers 1..M-5 in cidx
    static_assert(LB * (M - 5) <= 32, "slot fields do not fit one word");
    __host__ __device__ static constexpr unsigned rep_field() {   // 1 at the lowest bit of every field
        unsigned r = 0;
        for (int i = 0; i < M - 5; ++i) r |= 1u << (LB * i);
        return r;
    }
    static_assert((FRAME_WORDS * 4) % 16 == 0 && (OFF_H * 4) % 16 == 0 && (OFF_KEYS * 4) % 16 == 0, "alignment");
    // PCB_Q_MW: warps of the one block of an SM (227 KB minus the reserved 1 KB; 18 warps keep 113 registers)
    static constexpr int WMAX = (232448 - 1024) / TOTAL < PCB_Q_WCAP ? (232448 - 1024) / TOTAL : PCB_Q_WCAP;
    static constexpr bool MW = LL >= 16 ? PCB_Q_MW_BIG : PCB_Q_MW;
};

// f for frames whose channel values are all 0 or within [2^-40, 2^100] (see decode_scl8.cu):
// min(|a|,|b|) * (a*b < 0 ? -1 : 1) (SCLDecoderMinsum.cpp:40-42) equals min(|a|,|b|) with the
// XOR of the two sign bits there (the product neither underflows nor overflows), which is one
// instruction: FMNMX.XORSIGN |a|, |b|.  A zero result may carry a minus sign where the reference
// has +0; no decision, metric update or later f/g distinguishes the two.
__device__ __forceinline__ float ff(float a, float b) {
    float r;
    asm("min.xorsign.abs.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
// g (SCLDecoderMinsum.cpp:51-53) with the partial-sum bit already at the sign position
__device__ __forceinline__ float gs(float a, float b, uint32_t signbit) {
    return b + __uint_as_float(__float_as_uint(a) ^ signbit);
}
__device__ __forceinline__ uint32_t sgn(uint32_t w, int pos) { return (w << (31 - pos)) & 0x80000000u; }
__device__ __forceinline__ float4 ff4(float4 a, float4 b) {
    return make_float4(ff(a.x, b.x), ff(a.y, b.y), ff(a.z, b.z), ff(a.w, b.w));
}
// b + a as two packed adds (FADD2: one issue slot for two lanes of the float4)
__device__ __forceinline__ float4 add4(float4 b, float4 a) {
    unsigned long long a0, a1, b0, b1, r0, r1;
    asm("mov.b64 %0, {%1,%2};" : "=l"(a0) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(a1) : "f"(a.z), "f"(a.w));
    asm("mov.b64 %0, {%1,%2};" : "=l"(b0) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(b1) : "f"(b.z), "f"(b.w));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r0) : "l"(b0), "l"(a0));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r1) : "l"(b1), "l"(a1));
    float4 r;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(r0));
    asm("mov.b64 {%0,%1}, %2;" : "=f"(r.z), "=f"(r.w) : "l"(r1));
    return r;
}
// partial-sum bits of the 4 elements at positions pos .. pos+3 of cw (pos <= 28)
__device__ __forceinline__ float4 gs4(float4 a, float4 b, uint32_t cw, int pos) {
#if PCB_Q_FADD2
    const uint32_t t = cw << (28 - pos);       // bit of element 3 at the sign position
    a.x = __uint_as_float(__float_as_uint(a.x) ^ ((t << 3) & 0x80000000u));
    a.y = __uint_as_float(__float_as_uint(a.y) ^ ((t << 2) & 0x80000000u));
    a.z = __uint_as_float(__float_as_uint(a.z) ^ ((t << 1) & 0x80000000u));
    a.w = __uint_as_float(__float_as_uint(a.w) ^ (t & 0x80000000u));
    return add4(b, a);
#else
    return make_float4(gs(a.x, b.x, sgn(cw, pos)), gs(a.y, b.y, sgn(cw, pos + 1)), gs(a.z, b.z, sgn(cw, pos + 2)),
                       gs(a.w, b.w, sgn(cw, pos + 3)));
#endif
}

// max over the 8 lanes of a frame: four full-warp reductions (one uniform-datapath CREDUX each,
// ~49 cycles for the four) instead of a three-level shuffle butterfly (~110 cycles); a
// __reduce_*_sync with an 8-lane member mask compiles to a divergent software loop (~310 cycles).
template <int FPW>
__device__ __forceinline__ unsigned group_max(unsigned x, int fme) {
    if constexpr (FPW == 1) return __reduce_max_sync(FULL, x);
    if constexpr (FPW >= 8) {     // 2 or 4 lanes per frame: a shuffle butterfly beats 8 or 16 reductions
#pragma unroll
        for (int o = 16 / FPW; o; o >>= 1) x = max(x, __shfl_xor_sync(FULL, x, o));
        return x;
    }
    unsigned r = 0;
#pragma unroll
    for (int g = 0; g < FPW; g++) {
        const unsigned v = __reduce_max_sync(FULL, fme == g ? x : 0u);
        if (fme == g) r = v;
    }
    return r;
}

// Partial-sum combine into a layer of size D = 2^d2 >= 32 (SCLDecoderMinsum.cpp:287-305), one
// lane per path: T holds the last 32 bits; the destination layer m-d2 sits at heap words
// [D/32, 2D/32) of this slot; layer 0 (the final codeword estimate, written at the very last
// phase) goes to the then-dead LLR region instead.  Out of line: taken once per four subtrees.
template <int LB, bool GH>
__device__ __noinline__ uint32_t heap_combine(uint32_t* Pw, uint32_t* H, uint32_t* HG, int hs, int splitw, int p, uint32_t cidx, int d2, uint32_t T, int m) {
    constexpr int LL = 1 << LB;
    const int D = 1 << d2, NW = 1 << (m - 5);
    // word i of the destination layer's slot is dp[i * ds]: the then-dead LLR region for layer 0,
    // the global part of the heap ([word][slot]) for the big layers, shared memory otherwise
    uint32_t* dp;
    int ds = 1;
    if (d2 == m) dp = Pw + p * (NW + 1) - NW;
    else if (GH && D / 32 >= splitw) { dp = HG + p - splitw * LL; ds = LL; }
    else dp = H + p * hs;
    __syncwarp();
    dp[(2 * D / 32 - 1) * ds] = T;
#pragma unroll 1
    for (int w = 32; w < D; w <<= 1) {
        const int lgw = 31 - __clz(w), layc = m - lgw, nw = w >> 5;
        const int sc = (cidx >> (LB * (layc - 1))) & (LL - 1);
        const uint32_t* sp;
        int ss = 1;
        if (GH && nw >= splitw) { sp = HG + sc + (nw - splitw) * LL; ss = LL; }
        else sp = H + sc * hs + nw;
#pragma unroll 1
        for (int kk = 0; kk < nw; kk++)
            dp[((2 * D - 2 * w) / 32 + kk) * ds] = sp[kk * ss] ^ dp[((2 * D - w) / 32 + kk) * ds];
    }
    __syncwarp();
    if (d2 < m) {
        const int sh = LB * (m - d2 - 1);
        cidx = (cidx & ~(((1u << LB) - 1) << sh)) | ((uint32_t)p << sh);
    }
    return cidx;
}

// Per-path state a fork can replace.  r = u | take << 1 | source lane << 2.
struct PathState {
    float met;
    uint32_t pidx, cidx, w0, r;
    float4 r4;          // register LLR layers (a dead one is copied too: one code path for all phases)
    float r2_0, r2_1;
    uint32_t maxmet;    // largest metric of the frame: in before, out after the fork
};

// Shortcut of the median select for the commonest non-trivial outcome.  Per frame, B = the paths
// whose penalised candidate hi = met + |llr| is not above the largest metric maxmet.  If B is
// empty the frame takes its fast exit; if B = {q0}, hi(q0) < maxmet strictly and exactly one path
// p0 holds maxmet, then the 8 smallest of the 16 candidates are the 7 other unpenalised ones plus
// both candidates of q0, with a strict gap on either side of the cut (no tie for pass 2 to
// break): p0 is killed and its slot receives the clone of q0 (:372-389).  Returns false (and the
// caller runs the full select) unless every frame of the warp is in one of the two cases.
template <int LL>
__device__ __forceinline__ bool single_swap(const PathState& s, float A0, int lane, uint32_t& r) {
    constexpr unsigned GM = LL == 32 ? 0xffffffffu : (1u << LL) - 1;
    const int gsh = lane & (32 - LL);
    const uint32_t lo = __float_as_uint(s.met), hi = __float_as_uint(s.met + fabsf(A0));
    const unsigned gB = (__ballot_sync(FULL, !(hi > s.maxmet)) >> gsh) & GM;
    const unsigned gE = (__ballot_sync(FULL, lo == s.maxmet) >> gsh) & GM;
    const unsigned gS = (__ballot_sync(FULL, hi == s.maxmet) >> gsh) & GM;
    const bool cond = gB == 0u || (__popc(gB) == 1 && __popc(gE) == 1 && gS == 0u);
    if (!__all_sync(FULL, cond)) return false;
    const int q0 = gsh + __ffs(gB) - 1, p0 = gsh + __ffs(gE) - 1;
    if (gB != 0u && lane == p0) r = 3u | ((uint32_t)q0 << 2);                       // clone of q0, bit 1
    else r = ((gB != 0u && lane == q0) ? 0u : (A0 < 0.f ? 1u : 0u)) | ((uint32_t)lane << 2);   // q0 keeps bit 0
    return true;
}

// continuePaths_UnfrozenBit (SCLDecoderMinsum.cpp:315-400) without the fast exit, for the four
// frames of the warp at once (every step is confined to the 8-lane group of a frame).
template <int LL, bool MERGE>
__device__ __noinline__ PathState fork_slow(PathState s, float A0, int lane, int ninfo, uint32_t* keys) {
    constexpr int FPW = 32 / LL, LB = LL == 2 ? 1 : LL == 4 ? 2 : LL == 8 ? 3 : LL == 16 ? 4 : 5;
    constexpr unsigned GM = LL == 32 ? 0xffffffffu : (1u << LL) - 1;
    const int p = lane & (LL - 1), gsh = lane & (32 - LL);
    const float pf0 = s.met + fmaxf(-A0, 0.f);   // pen(llr,0) = |llr| iff llr < 0   (:308-313)
    const float pf1 = s.met + fmaxf(A0, 0.f);    // pen(llr,1) = |llr| iff llr > 0
    uint32_t r;
    if (ninfo < LB) {
        // growth (:329-341): live slots are [L-nact, L); parent q clones into slot 2*pbase-1-q
        const int nact = 1 << ninfo, pbase = LL - nact;
        const bool take = p >= pbase - nact && p < pbase;
        r = take ? (3u | ((uint32_t)(gsh + 2 * pbase - 1 - p) << 2)) : ((uint32_t)lane << 2);
    } else if (single_swap<LL>(s, A0, lane, r)) {
    } else {
        const uint32_t k0 = __float_as_uint(pf0), k1 = __float_as_uint(pf1);
        bool keep0, keep1;
        unsigned b0, b1;
#define PCB_LT(c, x, k) asm("{ .reg .u32 t; add.cc.u32 t, %1, %2; addc.u32 %0, %0, 0; }" : "+r"(c) : "r"(x), "r"(k));
        if constexpr (MERGE) {
        // The survivors are the L smallest of the 2L candidates in the order (value, candidate number
        // 2p + b) (:342-371: everything below the median, then ties at the median in candidate
        // order).  Every path has a smaller candidate lo (its metric) and a larger one hi; sort the
        // L lo's and the L hi's separately (ties inside a list fall in path order, which is the
        // candidate order) and merge: with a_0 < .. < a_L-1 the lo's and b_0 < .. the hi's, exactly
        // c = #{i : b_i < a_(L-1-i)} hi's survive (the predicate is monotone in i), together with
        // the L - c smallest lo's.  2 x L x L compares per frame instead of 2L x 2L.
        const bool lob = k1 < k0;                 // bit of the smaller candidate (bit 0 when equal)
        const uint32_t lov = lob ? k1 : k0, hiv = lob ? k0 : k1;
        keys[p] = ~lov;                           // complemented: key < kc <=> carry-out of (kc + ~key)
        keys[LL + p] = ~hiv;
        __syncwarp();
        int ca = 0, cb = 0, da = 0, db = 0;
#pragma unroll 4
        for (int q = 0; q < LL / 4; q++) {
            const uint4 ka = *reinterpret_cast<const uint4*>(keys + 4 * q);
            const uint4 kb = *reinterpret_cast<const uint4*>(keys + LL + 4 * q);
            PCB_LT(ca, ka.x, lov) PCB_LT(cb, kb.x, hiv) PCB_LT(da, ka.y, lov) PCB_LT(db, kb.y, hiv)
            PCB_LT(ca, ka.z, lov) PCB_LT(cb, kb.z, hiv) PCB_LT(da, ka.w, lov) PCB_LT(db, kb.w, hiv)
        }
        int ra = ca + da, rb = cb + db;           // ranks inside the two lists, so far without the ties
        // the strict ranks of a list are a permutation unless two of its values are equal (rare)
        if ((__reduce_or_sync(FULL, 1u << (gsh + ra)) & __reduce_or_sync(FULL, 1u << (gsh + rb))) != FULL) {
            const unsigned below = (1u << p) - 1;
            ra += __popc((__match_any_sync(FULL, lov) >> gsh) & GM & below);
            rb += __popc((__match_any_sync(FULL, hiv) >> gsh) & GM & below);
        }
        __syncwarp();                             // the keys are consumed
        uint2* SA = reinterpret_cast<uint2*>(keys);
        uint2* SB = SA + LL;
        SA[LL - 1 - ra] = make_uint2(lov, 2u * p + (lob ? 1u : 0u));      // lo's, descending
        SB[rb] = make_uint2(hiv, 2u * p + (lob ? 0u : 1u));               // hi's, ascending
        __syncwarp();
        const uint2 av = SA[p], bv = SB[p];
        const bool hi_first = bv.x < av.x || (bv.x == av.x && bv.y < av.y);
        const int c = __popc((__ballot_sync(FULL, hi_first) >> gsh) & GM);
        const bool keep_lo = ra < LL - c, keep_hi = rb < c;
        keep0 = lob ? keep_hi : keep_lo;
        keep1 = lob ? keep_lo : keep_hi;
        b0 = (__ballot_sync(FULL, keep0) >> gsh) & GM;
        b1 = (__ballot_sync(FULL, keep1) >> gsh) & GM;
        } else {
        // stable rank of both candidates among the 2L of the frame; keys are stored complemented:
        // key < kc  <=>  carry-out of (kc + ~key)
        *reinterpret_cast<uint2*>(keys + 2 * p) = make_uint2(~k0, ~k1);
        __syncwarp();
        int c0 = 0, c1 = 0, d0 = 0, d1 = 0;   // two accumulators per candidate: shorter dependent chains
#pragma unroll 4
        for (int q = 0; q < LL / 2; q++) {
            const uint4 kk = *reinterpret_cast<const uint4*>(keys + 4 * q);
            PCB_LT(c0, kk.x, k0) PCB_LT(c1, kk.x, k1) PCB_LT(d0, kk.y, k0) PCB_LT(d1, kk.y, k1)
            PCB_LT(c0, kk.z, k0) PCB_LT(c1, kk.z, k1) PCB_LT(d0, kk.w, k0) PCB_LT(d1, kk.w, k1)
        }
        keep0 = c0 + d0 < LL; keep1 = c1 + d1 < LL;
        b0 = (__ballot_sync(FULL, keep0) >> gsh) & GM; b1 = (__ballot_sync(FULL, keep1) >> gsh) & GM;
        const bool tie = __popc(b0) + __popc(b1) != LL;        // a tie straddles the median (:362-371)
        if (__any_sync(FULL, tie)) {
            const unsigned top = max(keep0 ? k0 : 0u, keep1 ? k1 : 0u);
            const unsigned tkey = group_max<FPW>(top, lane / LL);
            const unsigned t0 = (__ballot_sync(FULL, keep0 && k0 == tkey) >> gsh) & GM;
            const unsigned t1 = (__ballot_sync(FULL, keep1 && k1 == tkey) >> gsh) & GM;
            const int need = LL - (__popc(b0) + __popc(b1) - __popc(t0) - __popc(t1));
            const unsigned lt = (1u << p) - 1;
            const int before0 = __popc(t0 & lt) + __popc(t1 & lt), before1 = before0 + ((t0 >> p) & 1);
            if (tie) {
                keep0 = keep0 && (k0 < tkey || before0 < need);
                keep1 = keep1 && (k1 < tkey || before1 < need);
            }
            b0 = (__ballot_sync(FULL, keep0) >> gsh) & GM;
            b1 = (__ballot_sync(FULL, keep1) >> gsh) & GM;
        }
        }
#undef PCB_LT
        const unsigned nonem = ~(b0 | b1) & GM, bothm = b0 & b1;
        if ((nonem >> p) & 1) {
            // killed slots are pushed ascending (:372-376) and popped by the cloning parents in
            // ascending order (:382-389): the j-th killed slot receives the clone of the j-th
            // HIGHEST parent
            const int j = __popc(nonem & ((1u << p) - 1));
            unsigned bm = bothm;
            if constexpr (LL <= 8) {       // at most L/2 slots are killed
                if (j > 0) bm &= ~(0x80000000u >> __clz(bm));
                if (j > 1) bm &= ~(0x80000000u >> __clz(bm));
                if (j > 2) bm &= ~(0x80000000u >> __clz(bm));
            } else {
#pragma unroll 1
                for (int jj = 0; jj < j; jj++) bm &= ~(0x80000000u >> __clz(bm));
            }
            r = 3u | ((uint32_t)(gsh + 31 - __clz(bm)) << 2);
        } else {
            r = (keep0 ? 0u : 1u) | ((uint32_t)lane << 2);
        }
    }
    s.met = (r & 1) ? pf1 : pf0;
    // clones copy the parent's state (:145-156 without copying arrays)
    __syncwarp();   // orders the parents' shared-memory writes before the clones' lazy reads; keys reusable
    const int sl = (r >> 2) & 31;
    const float m1 = __shfl_sync(FULL, pf1, sl);
    const uint32_t pi = __shfl_sync(FULL, s.pidx, sl);
    const uint32_t ci = __shfl_sync(FULL, s.cidx, sl);
    const uint32_t ww = __shfl_sync(FULL, s.w0, sl);
    const float x4 = __shfl_sync(FULL, s.r4.x, sl), y4 = __shfl_sync(FULL, s.r4.y, sl);
    const float z4 = __shfl_sync(FULL, s.r4.z, sl), w4 = __shfl_sync(FULL, s.r4.w, sl);
    const float a2 = __shfl_sync(FULL, s.r2_0, sl), b2 = __shfl_sync(FULL, s.r2_1, sl);
    if (r & 2u) {
        s.met = m1; s.pidx = pi; s.cidx = ci; s.w0 = ww;
        s.r4 = make_float4(x4, y4, z4, w4); s.r2_0 = a2; s.r2_1 = b2;
    }
    s.r = r;
    s.maxmet = group_max<FPW>(__float_as_uint(s.met), lane / LL);
    return s;
}

template <int M, int LL, typename YT, bool ROWS>
struct Dec {
    using G = Geo<M, LL>;
    static constexpr int FPW = G::FPW, LB = G::LB;
    int lane, p;
    // channel values of the frames of this quad: ROWS: the block's own float rows (frame f of the quad at
    // ysrc + f * N, written by the quad prologue: narrowed doubles or generated frames), else the batch itself
    const YT* ysrc;
    __device__ __forceinline__ const YT* yrow(int f) const {
        if constexpr (ROWS) return ysrc + f * G::N;
        else return ysrc + (size_t)min(fbase + f, Bm1) * G::N;
    }
    int fbase, Bm1;    // first frame of the quad, last frame of the batch (tail quads repeat it)
    float* P;          // this frame's LLR layers
    uint32_t* H;       // this frame's partial-sum heaps
    uint32_t* HG;      // GH: their big layers in global memory, word w of slot s at HG[(w - SPLITW) * LL + s]
    uint32_t* keys;    // this frame's 16 candidate keys
    float* W;          // the warp's whole slice (4 frames)
    float4* L3;        // GL3: this frame's layer-3 arrays in global memory, element (k, slot) at L3[k*8 + slot]
    float4* L4;        // GL4: same for layer 4
    float4* L5;        // GL5: same for layer 5
    float met;
    uint32_t maxmet;   // bit pattern of the largest metric of this frame (valid whenever the next phase is an information bit)
    uint32_t pidx, cidx, w0;
    float4 vA, vB;     // layer LBOT (size 8) of the current subtree, handed from upper() to leaf8()
    float4 r4;         // size-4 layer
    float r2_0, r2_1;  // size-2 layer
    float A0;          // leaf LLR
    int ninfo;         // information bits seen so far (warp-uniform): the list is full from the 4th on

    __device__ __forceinline__ float* arr(int lam, int slot) const { return P + G::off(lam) + slot * G::stride(lam); }
    __device__ __forceinline__ int pfield(int lam) const { return (pidx >> (LB * (lam - G::LTOP))) & (LL - 1); }
    __device__ __forceinline__ int cfield(int lam) const { return (cidx >> (LB * (lam - 1))) & (LL - 1); }
    // word w32 (32 elements each) of partial-sum layer `lam` (size >= 32), elements from `base`
    __device__ __forceinline__ uint32_t cword(int lam, int base, int w32) const {
        const int w = (G::size(lam) >> 5) + (base >> 5) + w32;
        if (G::GH && (G::size(lam) >> 5) >= G::SPLITW) return HG[(w - G::SPLITW) * LL + cfield(lam)];
        return H[cfield(lam) * G::HS + w];
    }

    // ---- layer 3 straight from the channel values (layers 0, 1, 2 are virtual) ---------
    // ev = i0 / (N/8) in 0..7: bit 2 -> layer 1 is a g-layer, bit 1 -> layer 2, bit 0 -> layer 3.
    // L1[x] = op1(y[x], y[x+4E]), L2[x] = op2(L1[x], L1[x+2E]), L3[j] = op3(L2[j], L2[j+E]), E = N/8.
    __device__ __forceinline__ float4 load_y4(const YT* y, int q) const {
        if (sizeof(YT) == 8) {
            const double2 a = *reinterpret_cast<const double2*>(y + 4 * q);
            const double2 b = *reinterpret_cast<const double2*>(y + 4 * q + 2);
            return make_float4((float)a.x, (float)a.y, (float)b.x, (float)b.y);   // :406-408
        } else {
            return *reinterpret_cast<const float4*>(y + 4 * q);
        }
    }
    // GL3: layers 3 and 4 from the channel values.  Layer 3 goes to the global scratch, layer 4
    // (always an f-layer here) to shared memory.  With a g-layer 1 the channel values are staged
    // in rounds of 8+8 float4 columns (k and k + Q3/2, the two operands of a layer-4 output), so
    // that the staging area never overlaps the layer-4 arrays being written.
    // Stage layer 1 = f(y[x], y[x + N/2]) (the same for every path) of the four frames.
    __device__ __forceinline__ void stage_l1f(const YT* ybase) {
        constexpr int N4 = G::N / 4, SU = PCB_Q_STU;
#pragma unroll 1
        for (int f = 0; f < FPW; f++) {
            const YT* y = yrow(f);
            float* Rf = W + f * G::FRAME_WORDS + G::STAGE_OFF;
#pragma unroll SU
            for (int q = lane; q < N4 / 2; q += 32)
                *reinterpret_cast<float4*>(Rf + 4 * q) = ff4(load_y4(y, q), load_y4(y, q + N4 / 2));
        }
    }

    // chunk c of the channel values of the four frames (columns 4c .. 4c+3 of each half, 8 rows):
    // 64 float4 per frame, one cp.async group (an empty one behind the last chunk)
    __device__ __forceinline__ void stage_chunk(const YT* ybase, int c) {
        constexpr int Q = G::Q3, CU = PCB_Q_CPU;
        static_assert(sizeof(YT) == 4, "cp.async staging copies float channel values");
        if (c < Q / 8) {
#pragma unroll CU
            for (int i = 0; i < 2 * FPW; i++) {    // unit v = i*32 + lane: frame i/2, then t (8), h (2), k4 (4)
                const int v = i * 32 + lane, w = v & 63, t = w >> 3, h = (w >> 2) & 1, kk = (c & 1) * 4 + (w & 3);
                const YT* y = yrow((i >> 1));
                const unsigned sa = (unsigned)__cvta_generic_to_shared(W + (i >> 1) * G::FRAME_WORDS + G::STAGE_OFF + ((t * 2 + h) * 8 + kk) * 4);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(y + 4 * (t * Q + h * (Q / 2) + (c >> 1) * 8 + kk)) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

    __device__ __forceinline__ void top_gl3(int ev, const YT* ybase, bool g4 = false) {
        constexpr int E = G::N / 8, N4 = G::N / 4, Q = G::Q3, NR = Q / 16, TU = PCB_Q_TOPU;
        [[maybe_unused]] constexpr int SG = PCB_Q_STG;
        const bool g1 = ev & 4, g2 = ev & 2, g3 = ev & 1;
        const float* R = P + G::STAGE_OFF;
        float* dst4 = G::GL4 ? nullptr : arr(4, p);
        [[maybe_unused]] float4* dst3 = G::RL3 ? nullptr : L3 + p;
        __syncwarp();                            // everyone is done with the previous subtree
        const bool chunked = g1 || G::CHUNKF;    // channel values staged chunk by chunk (else: layer 1 = f staged whole)
        if (!chunked) {
            stage_l1f(ybase);
            __syncwarp();
        }
#if PCB_Q_CPA
        // g-layer 1: the channel values go straight from L2 to the staging area with cp.async, in
        // chunks of 4 columns x 2 halves x 8 rows per frame (chunk c = columns c*4 .. c*4+3); chunk
        // c+2 is in flight while chunk c is consumed, so the L2 latency hides behind the arithmetic.
        if (chunked) {
#pragma unroll 1
            for (int c = 0; c < 2; c++) stage_chunk(ybase, c);
        }
#endif
#pragma unroll 1
        for (int r = 0; r < NR; r++) {
#if !PCB_Q_CPA
            if (g1) {
                if (r) __syncwarp();
#pragma unroll SG
                for (int i = 0; i < 4 * FPW; i++) {   // unit u = i*32 + lane: frame i/4, then t (8), h (2), kk (8)
                    const int u = i * 32 + lane, t = (u >> 4) & 7, h = (u >> 3) & 1, kk = u & 7;
                    const YT* y = yrow((i >> 2));
                    *reinterpret_cast<float4*>(W + (i >> 2) * G::FRAME_WORDS + G::STAGE_OFF + (u & 127) * 4) =
                        load_y4(y, t * Q + h * (Q / 2) + r * 8 + kk);
                }
                __syncwarp();
            }
#endif
            uint32_t c1[4][2], c2[2][2], c3[2];
            [[maybe_unused]] uint32_t c4 = 0;
            if constexpr (G::RL3) { if (g4) c4 = cword(4, 0, r); }
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int w32 = h * (Q / 16) + r;
                c3[h] = g3 ? cword(3, 0, w32) : 0u;
#pragma unroll
                for (int t = 0; t < 4; t++) c1[t][h] = g1 ? cword(1, t * E, w32) : 0u;
                c2[0][h] = g2 ? cword(2, 0, w32) : 0u; c2[1][h] = g2 ? cword(2, E, w32) : 0u;
            }
#if PCB_Q_CPA
#pragma unroll 1
            for (int hf = 0; hf < 2; hf++) {
            if (chunked) {
                asm volatile("cp.async.wait_group 1;" ::: "memory");
                __syncwarp();
            }
#pragma unroll TU
            for (int kk = 4 * hf; kk < 4 * hf + 4; kk++) {
#else
#pragma unroll TU
            for (int kk = 0; kk < 8; kk++) {
#endif
                float4 o[2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int k = h * (Q / 2) + r * 8 + kk;
                    float4 l1[4];
                    if (chunked) {
                        float4 yv[8];
#pragma unroll
                        for (int t = 0; t < 8; t++) yv[t] = *reinterpret_cast<const float4*>(R + ((t * 2 + h) * 8 + kk) * 4);
                        if (g1) {
#pragma unroll
                            for (int t = 0; t < 4; t++) l1[t] = gs4(yv[t], yv[t + 4], c1[t][h], 4 * kk);
                        } else if constexpr (G::CHUNKF) {
#pragma unroll
                            for (int t = 0; t < 4; t++) l1[t] = ff4(yv[t], yv[t + 4]);
                        }
                    } else {
#pragma unroll
                        for (int t = 0; t < 4; t++) l1[t] = *reinterpret_cast<const float4*>(R + t * E + 4 * k);
                    }
                    float4 l2[2];
                    if (g2) { l2[0] = gs4(l1[0], l1[2], c2[0][h], 4 * kk); l2[1] = gs4(l1[1], l1[3], c2[1][h], 4 * kk); }
                    else { l2[0] = ff4(l1[0], l1[2]); l2[1] = ff4(l1[1], l1[3]); }
                    o[h] = g3 ? gs4(l2[0], l2[1], c3[h], 4 * kk) : ff4(l2[0], l2[1]);
                    if constexpr (!G::RL3) dst3[k * LL] = o[h];
                }
                if constexpr (G::RL3) L4[(r * 8 + kk) * LL + p] = g4 ? gs4(o[0], o[1], c4, 4 * kk) : ff4(o[0], o[1]);
                else if constexpr (G::GL4) L4[(r * 8 + kk) * LL + p] = ff4(o[0], o[1]);
                else *reinterpret_cast<float4*>(dst4 + 4 * (r * 8 + kk)) = ff4(o[0], o[1]);
            }
#if PCB_Q_CPA
            if (chunked) {
                __syncwarp();                    // the chunk is consumed: its slots take chunk c + 2
                stage_chunk(ybase, 2 * r + hf + 2);
            }
            }
#endif
        }
        __syncwarp();                            // the staging area is overwritten by layer 5 next
    }

    __device__ __forceinline__ void top3(int ev, const YT* ybase) {
        constexpr int E = G::N / 8, N4 = G::N / 4;
        const bool g1 = ev & 4, g2 = ev & 2, g3 = ev & 1;
        __syncwarp();                            // everyone is done with layers 4.. of the previous subtree
        if (g1) {
            // stage the channel values of the four frames as floats
#pragma unroll 1
            for (int f = 0; f < FPW; f++) {
                const YT* y = yrow(f);
                float* R = W + f * G::FRAME_WORDS + G::STAGE_OFF;
#pragma unroll 4
                for (int q = lane; q < N4; q += 32) *reinterpret_cast<float4*>(R + 4 * q) = load_y4(y, q);
            }
        } else {
            stage_l1f(ybase);
        }
        __syncwarp();
        const float* R = P + G::STAGE_OFF;
        float* dst = arr(3, p);
#pragma unroll 1
        for (int k0 = 0; k0 < E / 4; k0 += 8) {          // 8 float4 = 32 outputs per round
            uint32_t c1[4] = {0, 0, 0, 0}, c2[2] = {0, 0}, c3w = 0;
            if (g1) {
#pragma unroll
                for (int t = 0; t < 4; t++) c1[t] = cword(1, t * E, k0 >> 3);
            }
            if (g2) { c2[0] = cword(2, 0, k0 >> 3); c2[1] = cword(2, E, k0 >> 3); }
            if (g3) c3w = cword(3, 0, k0 >> 3);
#pragma unroll 2
            for (int kk = 0; kk < 8; kk++) {
                const int k = k0 + kk;
                float4 l1[4];
                if (g1) {
                    float4 yv[8];
#pragma unroll
                    for (int t = 0; t < 8; t++) yv[t] = *reinterpret_cast<const float4*>(R + t * E + 4 * k);
#pragma unroll
                    for (int t = 0; t < 4; t++) l1[t] = gs4(yv[t], yv[t + 4], c1[t], 4 * kk);
                } else {
#pragma unroll
                    for (int t = 0; t < 4; t++) l1[t] = *reinterpret_cast<const float4*>(R + t * E + 4 * k);
                }
                float4 l2[2], o;
                if (g2) { l2[0] = gs4(l1[0], l1[2], c2[0], 4 * kk); l2[1] = gs4(l1[1], l1[3], c2[1], 4 * kk); }
                else { l2[0] = ff4(l1[0], l1[2]); l2[1] = ff4(l1[1], l1[3]); }
                o = g3 ? gs4(l2[0], l2[1], c3w, 4 * kk) : ff4(l2[0], l2[1]);
                *reinterpret_cast<float4*>(dst + 4 * k) = o;
            }
        }
        __syncwarp();                            // staging area is overwritten by layer 4 next
    }

    // ---- one stored layer from the previous stored layer ------------------------------
    // The layer is a run-time value: one compact loop serves every size (the ten resident warps
    // of an SM sit at different places of the program, so code size is instruction-cache traffic).
    __device__ __forceinline__ int off_rt(int lam) const {      // == G::off(lam)
        return LL * ((G::N >> (G::LS - 1)) - (G::N >> (lam - 1))) + 4 * LL * (lam - G::LS);
    }
    // f-layer: own arrays
    __device__ __forceinline__ void layer_f(int lam) {
        const int S = G::N >> lam;
        const float* src = P + off_rt(lam - 1) + p * (2 * S + 4);
        float* dst = P + off_rt(lam) + p * (S + 4);
#pragma unroll 2
        for (int k = 0; k < S; k += 4)
            *reinterpret_cast<float4*>(dst + k) =
                ff4(*reinterpret_cast<const float4*>(src + k), *reinterpret_cast<const float4*>(src + S + k));
    }
    // g-layer (first layer of a chain): the source slot is the lazily shared one; partial sums of
    // sizes >= 32 come from the heap, smaller ones from w0 (bits [S, 2S))
    __device__ __forceinline__ void layer_g(int lam) {
        const int S = G::N >> lam;
        const float* src = P + off_rt(lam - 1) + pfield(lam - 1) * (2 * S + 4);
        float* dst = P + off_rt(lam) + p * (S + 4);
        const uint32_t* hw = H + cfield(lam) * G::HS + (S >> 5);
#pragma unroll 2
        for (int k = 0; k < S; k += 4) {
            const uint32_t cw = S >= 32 ? (hw[k >> 5] >> (k & 31)) : (w0 >> (S + k));
            *reinterpret_cast<float4*>(dst + k) =
                gs4(*reinterpret_cast<const float4*>(src + k), *reinterpret_cast<const float4*>(src + S + k), cw, 0);
        }
    }
    // ---- GL5: layers 3, 4, 5 in the global scratch -----------------------------------------
    __device__ __forceinline__ float4* glayer(int lam) const { return lam == 3 ? L3 : lam == 4 ? L4 : L5; }
    // Layer LAM (global) from the global layer LAM-1: g from the lazily shared source slot, or f
    // from this slot's own arrays.  Layer LAM+1 = f(layer LAM) comes next for this slot, so columns
    // j and j + S/8 are produced together and combined on the spot; it goes to the scratch when it is
    // a global layer too, else to shared memory.
    template <int LAM>
    __device__ __forceinline__ void global_fused(bool isg) {
        constexpr int S = G::size(LAM);
        const float4* src = glayer(LAM - 1) + (isg ? pfield(LAM - 1) : p);
        float4* dstg = glayer(LAM);
        float* dsts = LAM + 1 >= G::LS ? arr(LAM + 1 >= G::LS ? LAM + 1 : G::LS, p) : nullptr;
#pragma unroll 1
        for (int j0 = 0; j0 < S / 8; j0 += 2) {       // 8 loads in flight
            uint32_t cw0 = 0, cw1 = 0;
            if (isg) { cw0 = cword(LAM, 0, j0 >> 3) >> (4 * (j0 & 7)); cw1 = cword(LAM, 0, (j0 + S / 8) >> 3) >> (4 * (j0 & 7)); }
            float4 a[2][2], b[2][2];
#pragma unroll
            for (int i = 0; i < 2; i++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++) {
                    const int c = j0 + i + hh * (S / 8);
                    a[i][hh] = src[c * LL]; b[i][hh] = src[(c + S / 4) * LL];
                }
#pragma unroll
            for (int i = 0; i < 2; i++) {
                float4 lo, hi;
                if (isg) { lo = gs4(a[i][0], b[i][0], cw0, 4 * i); hi = gs4(a[i][1], b[i][1], cw1, 4 * i); }
                else { lo = ff4(a[i][0], b[i][0]); hi = ff4(a[i][1], b[i][1]); }
                dstg[(j0 + i) * LL + p] = lo;
                dstg[(j0 + i + S / 8) * LL + p] = hi;
                if constexpr (LAM + 1 >= G::LS) *reinterpret_cast<float4*>(dsts + 4 * (j0 + i)) = ff4(lo, hi);
                else glayer(LAM + 1)[(j0 + i) * LL + p] = ff4(lo, hi);
            }
        }
    }
    // shared-memory layer LAM from the global layer LAM-1
    template <int LAM>
    __device__ __forceinline__ void from_global(bool isg) {
        constexpr int S = G::size(LAM), NL = 4;      // (8 columns per iteration: -1.6 % at L = 32)
        const float4* src = glayer(LAM - 1) + (isg ? pfield(LAM - 1) : p);
        float* dst = arr(LAM, p);
#pragma unroll 1
        for (int k0 = 0; k0 < S / 4; k0 += NL) {
            const uint32_t cw = isg ? cword(LAM, 0, k0 >> 3) >> (4 * (k0 & 7)) : 0u;
            float4 a[NL], b[NL];
#pragma unroll
            for (int kk = 0; kk < NL; kk++) { a[kk] = src[(k0 + kk) * LL]; b[kk] = src[(k0 + kk + S / 4) * LL]; }
            if (isg) {
#pragma unroll
                for (int kk = 0; kk < NL; kk++) *reinterpret_cast<float4*>(dst + 4 * (k0 + kk)) = gs4(a[kk], b[kk], cw, 4 * kk);
            } else {
#pragma unroll
                for (int kk = 0; kk < NL; kk++) *reinterpret_cast<float4*>(dst + 4 * (k0 + kk)) = ff4(a[kk], b[kk]);
            }
        }
    }

    // GL3: g at layer 4 from the layer-3 arrays in global memory (L2)
    __device__ __forceinline__ void layer4_g_global() {
        constexpr int S = G::size(4), NJ = PCB_Q_L4N;
        const float4* src = L3 + pfield(3);
        if constexpr (G::GL4) {
            // layer 4 goes back to the scratch and, as layer 5 = f(layer 4) comes next for this
            // slot, columns j and j + S/8 are produced together and combined on the spot
            float* dst5 = arr(5, p);
#pragma unroll 1
            for (int j0 = 0; j0 < S / 8; j0 += NJ) {      // 4 * NJ loads in flight
                const uint32_t cw0 = cword(4, 0, j0 >> 3) >> (4 * (j0 & 7)), cw1 = cword(4, 0, (j0 + S / 8) >> 3) >> (4 * (j0 & 7));
                float4 a[NJ][2], b[NJ][2];
#pragma unroll
                for (int i = 0; i < NJ; i++)
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        const int c = j0 + i + hh * (S / 8);
                        a[i][hh] = src[c * LL]; b[i][hh] = src[(c + S / 4) * LL];
                    }
#pragma unroll
                for (int i = 0; i < NJ; i++) {
                    const float4 lo = gs4(a[i][0], b[i][0], cw0, 4 * i), hi = gs4(a[i][1], b[i][1], cw1, 4 * i);
                    L4[(j0 + i) * LL + p] = lo;
                    L4[(j0 + i + S / 8) * LL + p] = hi;
                    *reinterpret_cast<float4*>(dst5 + 4 * (j0 + i)) = ff4(lo, hi);
                }
            }
        } else {
            float* dst = arr(4, p);
#pragma unroll 1
            for (int k0 = 0; k0 < S / 4; k0 += 4) {          // 8 loads in flight; compact code
                const uint32_t cw = cword(4, 0, k0 >> 3) >> (4 * (k0 & 7));
                float4 a[4], b[4];
#pragma unroll
                for (int kk = 0; kk < 4; kk++) { a[kk] = src[(k0 + kk) * LL]; b[kk] = src[(k0 + kk + S / 4) * LL]; }
#pragma unroll
                for (int kk = 0; kk < 4; kk++)
                    *reinterpret_cast<float4*>(dst + 4 * (k0 + kk)) = gs4(a[kk], b[kk], cw, 4 * kk);
            }
        }
    }
    // GL4: layer 5 (shared memory) from the layer-4 arrays in global memory
    __device__ __forceinline__ void layer5_global(bool isg) {
        constexpr int S = G::size(5), NL = PCB_Q_L5N < S / 4 ? PCB_Q_L5N : S / 4;
        const float4* src = L4 + (isg ? pfield(4) : p);
        float* dst = arr(5, p);
#pragma unroll 1
        for (int k0 = 0; k0 < S / 4; k0 += NL) {
            const uint32_t cw = isg ? cword(5, 0, k0 >> 3) >> (4 * (k0 & 7)) : 0u;
            float4 a[NL], b[NL];
#pragma unroll
            for (int kk = 0; kk < NL; kk++) { a[kk] = src[(k0 + kk) * LL]; b[kk] = src[(k0 + kk + S / 4) * LL]; }
            if (isg) {
#pragma unroll
                for (int kk = 0; kk < NL; kk++) *reinterpret_cast<float4*>(dst + 4 * (k0 + kk)) = gs4(a[kk], b[kk], cw, 4 * kk);
            } else {
#pragma unroll
                for (int kk = 0; kk < NL; kk++) *reinterpret_cast<float4*>(dst + 4 * (k0 + kk)) = ff4(a[kk], b[kk]);
            }
        }
    }

    // ---- register chains: layers of size <= 32 --------------------------------------
    // From the first layer with at most 8 float4 (LR) down to LBOT a chain is computed in
    // registers: every layer is stored once (a g reads it later, maybe from a clone) but never
    // re-read, and layer LBOT is handed to the leaf in vA/vB.
    static constexpr int LR = (G::LBOT - PCB_Q_CHAIN > G::LS) ? G::LBOT - PCB_Q_CHAIN : G::LS + 1;  // size(LR) = 8 << PCB_Q_CHAIN
    template <int LAM, int NF4>
    __device__ __forceinline__ void f_down(float4 (&v)[NF4]) {
        float* dst = arr(LAM, p);
#pragma unroll
        for (int k = 0; k < NF4; k++) *reinterpret_cast<float4*>(dst + 4 * k) = v[k];
        if constexpr (LAM < G::LBOT) {
            float4 w[NF4 / 2];
#pragma unroll
            for (int k = 0; k < NF4 / 2; k++) w[k] = ff4(v[k], v[k + NF4 / 2]);
            f_down<LAM + 1, NF4 / 2>(w);
        } else {
            vA = v[0]; vB = v[1];
        }
    }
    // g at layer LAM (LR..LBOT) from the lazily shared layer LAM-1 in shared memory, then f down
    template <int LAM>
    __device__ __forceinline__ void chain_g() {
        constexpr int S = G::size(LAM), NF4 = S / 4;
        const float* src = arr(LAM - 1, pfield(LAM - 1));
        uint32_t cw;
        if constexpr (S >= 32) cw = cword(LAM, 0, 0);
        else cw = w0 >> S;
        float4 v[NF4];
#pragma unroll
        for (int k = 0; k < NF4; k++)
            v[k] = gs4(*reinterpret_cast<const float4*>(src + 4 * k), *reinterpret_cast<const float4*>(src + S + 4 * k), cw, 4 * k);
        f_down<LAM, NF4>(v);
    }
    // f at layer LR from this slot's own layer LR-1 in shared memory, then f down
    __device__ __forceinline__ void chain_f() {
        constexpr int S = G::size(LR), NF4 = S / 4;
        const float* src = arr(LR - 1, p);
        float4 v[NF4];
#pragma unroll
        for (int k = 0; k < NF4; k++)
            v[k] = ff4(*reinterpret_cast<const float4*>(src + 4 * k), *reinterpret_cast<const float4*>(src + S + 4 * k));
        f_down<LR, NF4>(v);
    }
    template <int LAM>
    __device__ __forceinline__ void chain_g_dispatch(int ll) {
        if constexpr (LAM <= G::LBOT) {
            if (ll == LAM) chain_g<LAM>();
            else chain_g_dispatch<LAM + 1>(ll);
        }
    }

    // LLR work at the first phase i0 = 8*blk of a size-8 subtree, down to layer LBOT
    __device__ __forceinline__ void upper(int blk, const YT* ybase) {
        static_assert(LR > G::LS, "register chains start below the first shared-memory layer");
        const int z = blk ? (__ffs(blk) - 1) : 31;
        int ll, own;    // layers own.. are recomputed, the f-layers below ll are still to do
        const bool topev = z >= M - 6;                     // i0 is a multiple of N/8
        if (topev || (G::RL3 && G::LBOT - z == 4)) {       // RL3: also at the g of layer 4 (layer 3 is recomputed, not read)
            if constexpr (G::GL3) { top_gl3(blk >> (M - 6), ybase, !topev); ll = 4; }
            else { top3(blk >> (M - 6), ybase); ll = 3; }
            own = topev ? G::LTOP : 4;
        } else {
            __syncwarp();
            ll = own = G::LBOT - z;
            if (ll >= LR) { chain_g_dispatch<LR>(ll); ll = G::LBOT; }
            else if (G::GL5 && ll == 4) { global_fused<4>(true); from_global<6>(false); ll = 6; }   // layers 4, 5, 6
            else if (G::GL5 && ll == 5) { global_fused<5>(true); ll = 6; }
            else if (G::GL5 && ll == 6) from_global<6>(true);
            else if (G::GL3 && ll == 4) { layer4_g_global(); if (G::GL4) ll = 5; }   // GL4: layer 5 done as well
            else if (G::GL4 && ll == 5) layer5_global(true);
            else layer_g(ll);
        }
        if (ll < G::LBOT) {
            int lam = ll + 1;
            if (G::GL5 && ll == 4) { global_fused<5>(false); lam = 7; }      // after the top stage
            else if (G::GL4 && ll == 4) { layer5_global(false); lam = 6; }
#pragma unroll 1
            for (; lam < LR; lam++) layer_f(lam);
            chain_f();
        }
        // layers >= own are now this slot's own
        const uint32_t low = (1u << (LB * (own - G::LTOP))) - 1;
        pidx = (pidx & low) | ((G::rep_field() * (uint32_t)p) & ~low);
    }

    // ---- fork / prune of one information bit (SCLDecoderMinsum.cpp:315-400) ----------
    __device__ __forceinline__ int fork() {
        if (ninfo >= LB) {
            // Fast exit (no kill, no clone): the candidates of path p are {met, met + |llr|}.  If
            // every penalised candidate of the frame is strictly above every unpenalised one, the
            // L smallest are exactly the L unpenalised ones (no tie possible).  Taken only when
            // all four frames of the warp agree.
            const float hi = met + fabsf(A0);
            if (__all_sync(FULL, __float_as_uint(hi) > maxmet)) return A0 < 0.f ? 1 : 0;   // max met < min hi
        }
        PathState s{met, pidx, cidx, w0, 0, r4, r2_0, r2_1, maxmet};
        s = fork_slow<LL, G::MERGE>(s, A0, lane, ninfo, keys);
        met = s.met; pidx = s.pidx; cidx = s.cidx; w0 = s.w0;
        r4 = s.r4; r2_0 = s.r2_0; r2_1 = s.r2_1; maxmet = s.maxmet;
        if (ninfo < LB) ninfo++;
        return s.r & 1;
    }

    // decision with the leaf LLR in A0
    __device__ __forceinline__ int decide(bool isinfo, bool nextinfo) {
        if (isinfo) return fork();
        met += fmaxf(-A0, 0.f);         // frozen value 0 (:413-418)
        // the fast exit of the next information bit compares against the largest metric of the frame
        if (nextinfo) maxmet = group_max<FPW>(__float_as_uint(met), lane / LL);
        return 0;
    }

    // partial-sum combine at the last phase of a subtree (t = 7): d2 = 3 + ctz(blk+1)
    // T: the size-4 partial sums of the last four phases
    __device__ __forceinline__ void setbit_last(int blk, uint32_t T) {
        const int d2 = min(M, 3 + __ffs(blk + 1) - 1);
        T = (((w0 >> 4) ^ T) & 15u) | (T << 4);
        if (d2 == 3) { w0 = (w0 & ~0xFF00u) | (T << 8); return; }
        T = (((w0 >> 8) ^ T) & 0xFFu) | (T << 8);
        if (d2 == 4) { w0 = (w0 & 0x0000FFFFu) | (T << 16); return; }
        T = (((w0 >> 16) ^ T) & 0xFFFFu) | (T << 16);
        cidx = heap_combine<LB, G::GH>(reinterpret_cast<uint32_t*>(P), H, HG, G::HS, G::SPLITW, p, cidx, d2, T, M);
    }

    // ---- the four phases of one size-4 subtree (its LLRs in r4, pairs (x,z) and (y,w)) --------
    // Three routines produce the same results as four successive phases of the reference loop
    // (SCLDecoderMinsum.cpp:409-431); each returns the size-4 partial sums [c0^d0, c1^d1, d0, d1].

    // all four positions frozen (value 0): no decisions, only the four metric updates in phase order
    __device__ __forceinline__ uint32_t rate0(bool nextinfo) {
        const float a = ff(r4.x, r4.z), b = ff(r4.y, r4.w);
        met += fmaxf(-ff(a, b), 0.f);
        met += fmaxf(-(b + a), 0.f);
        const float a2 = r4.z + r4.x, b2 = r4.w + r4.y;     // g with partial sums 0
        met += fmaxf(-ff(a2, b2), 0.f);
        met += fmaxf(-(b2 + a2), 0.f);
        if (nextinfo) maxmet = group_max<FPW>(__float_as_uint(met), lane / LL);
        return 0u;
    }

    // three frozen positions, then one information bit (the commonest mixed pattern)
    __device__ __forceinline__ uint32_t rep4() {
        const float a = ff(r4.x, r4.z), b = ff(r4.y, r4.w);
        met += fmaxf(-ff(a, b), 0.f);
        met += fmaxf(-(b + a), 0.f);
        r2_0 = r4.z + r4.x; r2_1 = r4.w + r4.y;             // g with partial sums 0
        met += fmaxf(-ff(r2_0, r2_1), 0.f);
        A0 = r2_1 + r2_0;                                   // g with u2 = 0
        maxmet = group_max<FPW>(__float_as_uint(met), lane / LL);
        return fork() ? 0xFu : 0u;                          // [u3, u3, u3, u3]
    }

    // Rate-1 node (all positions information bits) with the list full.  If every phase of the node
    // takes the fast exit, each path keeps its likelier bit at every phase and no metric changes, so
    // the node is an SC decode with hard decisions per path, and for those (min-sum, no zero LLR):
    //   * the node's partial sums are the signs of its input LLRs (u0 = sa^sb, u1 = sb for a pair;
    //     induction over the sizes), and
    //   * every g is a sum of two same-signed values (|g| = fl(|a|+|b|) >= max), every f a min, so
    //     the smallest leaf magnitude is the first leaf's, min_i |alpha_i|, exactly.
    // All fast exits of the node hold iff met + min_i |alpha_i| exceeds the largest metric of the
    // frame (fl(met + x) is monotone in x): ONE vote, no leaf LLR computed.  A zero input fails the
    // test (hi == met <= maxmet).  Nothing is committed on failure.
#if PCB_Q_R1FAST
    __device__ __forceinline__ static uint32_t signs4(float4 v) {
        return (__float_as_uint(v.x) >> 31) | ((__float_as_uint(v.y) >> 31) << 1) | ((__float_as_uint(v.z) >> 31) << 2) |
               ((__float_as_uint(v.w) >> 31) << 3);
    }
    __device__ __forceinline__ bool rate1(uint32_t& T4) {
        const float mn = fminf(fminf(fabsf(r4.x), fabsf(r4.y)), fminf(fabsf(r4.z), fabsf(r4.w)));
        const float hi = met + mn;
        if (!__all_sync(FULL, __float_as_uint(hi) > maxmet)) return false;
        T4 = signs4(r4);
        return true;
    }
    // the same for a whole size-8 subtree (its input LLRs in vA, vB): partial sums [sA ^ sB, sB]
    __device__ __forceinline__ bool rate1_8(uint32_t& T4) {
        const float ma = fminf(fminf(fabsf(vA.x), fabsf(vA.y)), fminf(fabsf(vA.z), fabsf(vA.w)));
        const float mb = fminf(fminf(fabsf(vB.x), fabsf(vB.y)), fminf(fabsf(vB.z), fabsf(vB.w)));
        const float hi = met + fminf(ma, mb);
        if (!__all_sync(FULL, __float_as_uint(hi) > maxmet)) return false;
        const uint32_t sA = signs4(vA), sB = signs4(vB);
        w0 = (w0 & ~0xF0u) | ((sA ^ sB) << 4);
        T4 = sB;
        return true;
    }
#else
    // all four positions information bits, list full: assume every phase takes the fast exit (each
    // path keeps its likelier bit, metrics unchanged), compute the four leaf LLRs lane-locally and
    // verify all four exits with ONE vote: min_t (met + |A_t|) = met + min_t |A_t| must exceed the
    // largest metric of the frame.  A zero LLR fails the test (hi == met), so below the test the
    // sign bit of A_t IS the decision (A_t < 0).  Nothing is committed on failure.
    __device__ __forceinline__ bool rate1(uint32_t& T4) {
        constexpr uint32_t SB = 0x80000000u;
        const float a = ff(r4.x, r4.z), b = ff(r4.y, r4.w);
        const float A0 = ff(a, b);
        const float A1 = b + __uint_as_float(__float_as_uint(a) ^ (__float_as_uint(A0) & SB));
        const uint32_t t01 = __float_as_uint(A0) ^ __float_as_uint(A1);            // sign = c0 = u0^u1
        const float a2 = r4.z + __uint_as_float(__float_as_uint(r4.x) ^ (t01 & SB));
        const float b2 = r4.w + __uint_as_float(__float_as_uint(r4.y) ^ (__float_as_uint(A1) & SB));
        const float A2 = ff(a2, b2);
        const float A3 = b2 + __uint_as_float(__float_as_uint(a2) ^ (__float_as_uint(A2) & SB));
        const float mn = fminf(fminf(fabsf(A0), fabsf(A1)), fminf(fabsf(A2), fabsf(A3)));
        const float hi = met + mn;
        if (!__all_sync(FULL, __float_as_uint(hi) > maxmet)) return false;
        const uint32_t t23 = __float_as_uint(A2) ^ __float_as_uint(A3);             // sign = d0 = u2^u3
        const uint32_t s1 = __float_as_uint(A1), s3 = __float_as_uint(A3);
        T4 = ((t01 ^ t23) >> 31) | (((s1 ^ s3) >> 31) << 1) | ((t23 >> 31) << 2) | ((s3 >> 31) << 3);
        return true;
    }
#endif

    // any pattern, phase by phase; ib: information flags of the four phases, bit 4 = next phase
    __device__ __forceinline__ uint32_t seq4(uint32_t ib) {
        r2_0 = ff(r4.x, r4.z); r2_1 = ff(r4.y, r4.w);
        A0 = ff(r2_0, r2_1);
        int u = decide(ib & 1, ib & 2);
        w0 = (w0 & ~2u) | ((uint32_t)u << 1);
        A0 = gs(r2_0, r2_1, sgn(w0, 1));
        u = decide(ib & 2, ib & 4);
        uint32_t T = (((w0 >> 1) ^ u) & 1u) | ((uint32_t)u << 1);            // size-2 partial sums
        w0 = (w0 & ~0xCu) | (T << 2);
        r2_0 = gs(r4.x, r4.z, sgn(w0, 2)); r2_1 = gs(r4.y, r4.w, sgn(w0, 3));
        A0 = ff(r2_0, r2_1);
        u = decide(ib & 4, ib & 8);
        w0 = (w0 & ~2u) | ((uint32_t)u << 1);
        A0 = gs(r2_0, r2_1, sgn(w0, 1));
        u = decide(ib & 8, ib & 16);
        T = (((w0 >> 1) ^ u) & 1u) | ((uint32_t)u << 1);
        return (((w0 >> 2) ^ T) & 3u) | (T << 2);                            // size 4
    }

    // the 8 phases of one size-8 subtree as two size-4 subtrees; layer LBOT is already computed.
    // ib: information flags of the 8 phases, bit 8 = first phase of the next subtree.
    __device__ __forceinline__ void leaf8(int blk, uint32_t ib) {
        uint32_t T4 = 0;
        int h = 0;
#if PCB_Q_R1FAST
        if ((ib & 0xFFu) == 0xFFu && ninfo >= LB && rate1_8(T4)) h = 2;
#endif
#pragma unroll 1
        for (; h < 2; h++, ib >>= 4) {
            if (h == 0) {                    // own arrays all the way down: layer LBOT comes in registers
                r4 = ff4(vA, vB);
            } else {                         // the size-8 layer may be a parent's
                __syncwarp();
                const float* v = arr(G::LBOT, pfield(G::LBOT));
                r4 = gs4(*reinterpret_cast<const float4*>(v), *reinterpret_cast<const float4*>(v + 4), w0, 4);
            }
            const uint32_t pat = ib & 15u;
            if (pat == 0u) T4 = rate0(ib & 16u);
            else if (pat == 8u) T4 = rep4();
            else if (pat == 15u && ninfo >= LB && rate1(T4)) {}
            else T4 = seq4(ib);
            if (h == 0) w0 = (w0 & ~0xF0u) | (T4 << 4);
        }
        setbit_last(blk, T4);
    }
};

}  // namespace q8

// Kernel parameters: the generic arguments plus the frozen pattern itself (warp-uniform control
// straight from the constant bank).
struct Scl8qArgs {
    DecodeArgs a;
    uint32_t info_w[65];   // n <= 2048, one zero word behind the last
    int* next_quad;        // work counter (zero at launch): quad gridDim.x + counter is the next one not yet taken
    float4* l3;            // n >= 1024: global scratch of the layer-3 arrays, Geo::L3_BLOCK_FLOAT4 per block
    uint32_t* xhat;        // out: [B][L][n/32] codeword estimates of every slot, the input of scl_select_kernel
    // Quad prologue (replaces the separate pre-pass): every frame is range-checked by the warp that decodes
    // it (frames outside the safe input range of the arithmetic go on the redo list of the generic kernel).
    // CONV kernels also narrow: frames below nf are floats at a.y, frame f >= nf is doubles at
    // y64 + (f - nf) * n (SCLDecoderMinsum.cpp:406-408 narrows every value first); the float rows of the
    // quad go to the block's slice of yscr (FPW * n floats, stays in L2) and the decoder reads them there.
    const double* y64;
    int nf;
    float* yscr;
    int check_range;       // 0: the producer of the batch has already range-checked it (on-device generator)
    // SRC = 2, the Monte-Carlo loop body in one kernel (PolarCodes.cpp:132-192): the warp GENERATES the frames of
    // its quad (payload, CRC, frozen bits, encoder, BPSK + AWGN: gen_common.cuh, the same functions and the same
    // Philox stream as generate_kernel) into its rows of yscr, decodes them, and counts frame / bit errors of
    // the k information bits against the payload it regenerates — no channel value ever reaches HBM.
    uint64_t seed, first_frame;
    float sigma;
    const uint32_t* gcrc;  // [16][IW] rows of the payload -> CRC map (crc = gcrc_c0 ^ XOR of the columns of the set bits)
    uint32_t gcrc_c0;
    uint32_t* skip;                    // bit f set: frame f is on the redo list, scl_select_kernel must not count it
    unsigned long long* mc_redo;       // [0] = count, then global frame ids of frames outside the safe input range
    int mc_redo_cap;
    int force_redo_mod;    // test hook: frames whose id is a multiple of this take the redo path
    int gen;               // 1: SRC = 2
};

namespace q8 {
// One frame of the Monte-Carlo loop body up to Transform (PolarCodes.cpp:133-158), made by the warp that will
// decode it: payload bits, CRC (GF(2)-affine in the payload: 16 parities per word), AddFrozenBits (frozen values
// 0), x = u F^(x)m, BPSK + AWGN into `row` (n floats, the block's own rows: they stay in L2).  Same device functions
// and the same Philox stream as generate_kernel.  SM: n/32 + (k+crc)/32 + 1 words of scratch (the frame's LLR region,
// dead between quads).  Returns false in the lanes that saw a value outside the safe input range of the kernel's arithmetic.
// Out of line and rolled on purpose: it runs once per frame, and every SASS row of this kernel competes for the
// instruction cache with the decoding loop of the 15 other warps of the SM.
template <int M>
__device__ __forceinline__ bool gen_frame(const Scl8qArgs& S, const Philox& ph, uint64_t fid, uint32_t* SM, float* row, int lane) {
    constexpr int n = 1 << M, NW = n / 32;
    const DecodeArgs& A = S.a;
    const int ninf = A.k + A.crc, IW = (ninf + 31) >> 5;
    const uint32_t f_lo = (uint32_t)fid, f_hi = (uint32_t)(fid >> 32);
    uint32_t* Wd = SM;            // u, then the codeword
    uint32_t* XI = SM + NW;       // payload + CRC
    // (loops with warp-uniform bounds and a lane predicate inside: the compiler then still knows that the warp is
    // converged everywhere in the kernel; lane-dependent trip counts here cost 40 % more SASS in the whole kernel)
#pragma unroll 1
    for (int w0 = 0; w0 < IW; w0 += 32) {                                                 // PolarCodes.cpp:133-135
        const int w = w0 + lane;
        const uint32_t v = payload_word(ph, f_lo, f_hi, min(w, IW - 1), A.k);
        if (w < IW) XI[w] = v;
    }
#pragma unroll 1
    for (int w0 = 0; w0 < NW; w0 += 32) { if (w0 + lane < NW) Wd[w0 + lane] = 0; }
    __syncwarp();
    if (A.crc) {                  // PolarCodes.cpp:136-155
        uint32_t acc = 0;
#pragma unroll 1
        for (int w0 = 0; w0 < IW; w0 += 32) {
            const int w = min(w0 + lane, IW - 1);
            const uint32_t x = w0 + lane < IW ? XI[w] : 0u;
            const uint4* g = reinterpret_cast<const uint4*>(S.gcrc + 16 * w);              // the 16 rows of word w
#pragma unroll 1
            for (int t4 = 0; t4 < 4; t4++) {
                const uint4 gv = __ldg(g + t4);
                acc ^= ((uint32_t)(__popc(x & gv.x) & 1) | ((uint32_t)(__popc(x & gv.y) & 1) << 1) | ((uint32_t)(__popc(x & gv.z) & 1) << 2) |
                        ((uint32_t)(__popc(x & gv.w) & 1) << 3)) << (4 * t4);
            }
        }
        const uint32_t r = __reduce_xor_sync(FULL, acc) ^ S.gcrc_c0;
        if (lane < A.crc && ((r >> (A.crc - 1 - lane)) & 1u)) atomicOr(&XI[(A.k + lane) >> 5], 1u << ((A.k + lane) & 31));   // MSB first
        __syncwarp();
    }
#pragma unroll 1
    for (int t0 = 0; t0 < ninf; t0 += 32) {                                                // AddFrozenBits, Encoder.cpp:17-33
        const int t = min(t0 + lane, ninf - 1);
        const int pos = A.info_pos[t];
        if (t0 + lane < ninf && ((XI[t >> 5] >> (t & 31)) & 1u)) atomicOr(&Wd[pos >> 5], 1u << (pos & 31));
    }
    __syncwarp();
    // x = u F^(x)m (Encoder.cpp:57-83): inside the words, then across them
#pragma unroll 1
    for (int w0 = 0; w0 < NW; w0 += 32) {
        const int w = w0 + lane;
        if (w < NW) {
            uint32_t x = Wd[w];
            x ^= (x >> 1) & 0x55555555u; x ^= (x >> 2) & 0x33333333u; x ^= (x >> 4) & 0x0F0F0F0Fu;
            x ^= (x >> 8) & 0x00FF00FFu; x ^= (x >> 16) & 0x0000FFFFu;
            Wd[w] = x;
        }
    }
    __syncwarp();
#pragma unroll 1
    for (int sw = 1; sw < NW; sw <<= 1) {
#pragma unroll 1
        for (int w0 = 0; w0 < NW; w0 += 32) {
            const int w = w0 + lane;
            if (w < NW && !(w & sw)) Wd[w] ^= Wd[w + sw];
        }
        __syncwarp();
    }
    // Transform (Encoder.cpp:35-41); the range check of the arithmetic rides along
    bool ok = true;
#pragma unroll 1
    for (int q0 = 0; q0 < n / 4; q0 += 32) {
        const int q = q0 + lane;
        const float4 v = channel4(ph, f_lo, f_hi, q, Wd[q >> 3] >> (4 * (q & 7)), S.sigma);
        reinterpret_cast<float4*>(row)[q] = v;
        const uint32_t ax = __float_as_uint(v.x) & 0x7fffffffu, ay = __float_as_uint(v.y) & 0x7fffffffu;
        const uint32_t az = __float_as_uint(v.z) & 0x7fffffffu, aw = __float_as_uint(v.w) & 0x7fffffffu;
        // |y| == 0 or 2^-40 (0x2B800000) <= |y| <= 2^100 (0x71800000)
        ok = ok && (ax == 0u || ax - 0x2B800000u <= 0x46000000u) && (ay == 0u || ay - 0x2B800000u <= 0x46000000u) &&
             (az == 0u || az - 0x2B800000u <= 0x46000000u) && (aw == 0u || aw - 0x2B800000u <= 0x46000000u);
    }
    if (S.force_redo_mod && (uint32_t)fid % (uint32_t)S.force_redo_mod == 0u) ok = false;      // (32-bit: a 64-bit modulo is 300 SASS rows)
    __syncwarp();
    return ok;                    // this lane's values; the caller votes (a vote in the kernel body keeps its control flow provably uniform)
}
}  // namespace q8

// One warp per block (blockDim.x == 32); a block decodes quads of frames.
template <int M, int LL, typename YT, int SRC>
__global__ void __launch_bounds__(q8::Geo<M, LL>::MW ? 32 * q8::Geo<M, LL>::WMAX : 32,
                                  q8::Geo<M, LL>::MW ? 1 : (LL != 8 ? 1 : q8::Geo<M, LL>::GL4 && M == 10 ? 16 : M == 10 ? 10 : 5))
    decode_scl8q_kernel(const __grid_constant__ Scl8qArgs S) {
    const DecodeArgs& A = S.a;
    using G = q8::Geo<M, LL>;
    constexpr int FPW = G::FPW;
    constexpr bool CONV = SRC == 1, GEN = SRC == 2, ROWS = SRC != 0;
    [[maybe_unused]] constexpr int PROU = PCB_Q_PROU;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr unsigned FULL = 0xffffffffu;
    // a warp is a virtual block (G::MW: several of them in one block)
    const int lane = G::MW ? threadIdx.x & 31 : threadIdx.x, wib = G::MW ? threadIdx.x >> 5 : 0, wpb = G::MW ? blockDim.x >> 5 : 1;
    const int vblock = blockIdx.x * wpb + wib, nvblocks = gridDim.x * wpb;
    unsigned char* const smem_warp = smem_raw + (size_t)wib * G::TOTAL;
    const int fme = lane / LL;
    q8::Dec<M, LL, YT, ROWS> D;
    D.lane = lane; D.p = lane & (LL - 1);
    D.W = reinterpret_cast<float*>(smem_warp);
    D.P = D.W + fme * G::FRAME_WORDS;
    D.H = reinterpret_cast<uint32_t*>(D.P) + G::OFF_H;
    D.HG = G::GH ? reinterpret_cast<uint32_t*>(S.l3 + (size_t)vblock * G::SCRATCH_BLOCK_FLOAT4 + G::L3_BLOCK_FLOAT4 + G::L4_BLOCK_FLOAT4 +
                                               G::L5_BLOCK_FLOAT4) + fme * ((G::NW - G::NW / 4) * LL)
                   : nullptr;
    D.keys = reinterpret_cast<uint32_t*>(D.P) + G::OFF_KEYS;
    D.L3 = G::GL3 ? S.l3 + (size_t)vblock * G::SCRATCH_BLOCK_FLOAT4 + fme * (G::Q3 * LL) : nullptr;
    D.L4 = G::GL4 ? S.l3 + (size_t)vblock * G::SCRATCH_BLOCK_FLOAT4 + G::L3_BLOCK_FLOAT4 + fme * (G::Q3 / 2 * LL) : nullptr;
    D.L5 = G::GL5 ? S.l3 + (size_t)vblock * G::SCRATCH_BLOCK_FLOAT4 + G::L3_BLOCK_FLOAT4 + G::L4_BLOCK_FLOAT4 + fme * (G::Q3 / 4 * LL) : nullptr;
    constexpr int n = G::N;
    const YT* ybase = reinterpret_cast<const YT*>(A.y);
    float* const yscr = ROWS ? S.yscr + (size_t)vblock * (FPW * G::N) : nullptr;
    D.ysrc = ROWS ? yscr : ybase;
    const int nquads = (A.B + FPW - 1) / FPW;
    [[maybe_unused]] const Philox ph{(uint32_t)S.seed, (uint32_t)(S.seed >> 32)};
    [[maybe_unused]] unsigned redo_mask = 0;                        // GEN: frames of the quad that scl_select_kernel must not count

    // The first quad of a block is its index, the following ones come from a work counter: decode
    // times differ from quad to quad, and a static stride leaves SMs idle at the end of the batch.
    int quad = vblock;
    while (quad < nquads) {
        D.fbase = FPW * quad; D.Bm1 = A.B - 1;
        // take the next quad now (an L2 prefetch of its channel values here, or late in the frame,
        // stopped paying once the top stage staged them with cp.async: 34.15 / 33.0 vs 34.23 M)
        int next_quad = 0;
#ifdef PCB_Q_STATIC
        next_quad = quad + nvblocks;
#else
        if (lane == 0) next_quad = nvblocks + atomicAdd(S.next_quad, 1);
        next_quad = __shfl_sync(FULL, next_quad, 0);
#endif
        D.met = 0.f; D.maxmet = 0; D.pidx = 0; D.cidx = 0; D.w0 = 0; D.ninfo = 0;
        D.r4 = make_float4(0.f, 0.f, 0.f, 0.f); D.r2_0 = D.r2_1 = D.A0 = 0.f;
        // ---- generating prologue: the Monte-Carlo loop body up to Transform (PolarCodes.cpp:133-158) ----
        if constexpr (GEN) {
            redo_mask = 0;
            __syncwarp();                              // the previous quad is done with the block's rows and the LLR region
#pragma unroll 1
            for (int f = 0; f < FPW; f++) {
                const uint64_t fid = S.first_frame + (uint64_t)min(FPW * quad + f, A.B - 1);
                const bool ok = q8::gen_frame<M>(S, ph, fid, reinterpret_cast<uint32_t*>(D.W + f * G::FRAME_WORDS), yscr + f * n, lane);
                if (!__all_sync(FULL, ok)) {
                    redo_mask |= 1u << f;
                    if (lane == 0 && FPW * quad + f < A.B) {
                        const unsigned long long slot = atomicAdd(S.mc_redo, 1ull);
                        if (slot < (unsigned long long)S.mc_redo_cap) S.mc_redo[1 + slot] = fid;
                    }
                }
            }
            __syncwarp();                              // the rows are written before anyone stages them
        }
        // ---- quad prologue: range check (and narrowing) of the quad's frames --------------------------
        if (!GEN && (CONV || S.check_range)) {
            if constexpr (CONV) __syncwarp();          // the previous quad is done with the block's rows
#pragma unroll 1
            for (int f = 0; f < FPW; f++) {
                const int frame = min(FPW * quad + f, A.B - 1);
                bool ok = true;
                auto chk = [&](float4 v) {
                    const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const uint32_t a = __float_as_uint(e[i]) & 0x7fffffffu;      // 2^-40 = 0x2B800000, 2^100 = 0x71800000
                        ok = ok && (a == 0u || (a - 0x2B800000u) <= (0x71800000u - 0x2B800000u));
                    }
                };
                if (CONV && frame >= S.nf) {
                    const double2* src = reinterpret_cast<const double2*>(S.y64 + (size_t)(frame - S.nf) * n);
                    float4* dst = reinterpret_cast<float4*>(yscr + f * n);
#pragma unroll PROU
                    for (int q0 = 0; q0 < n / 4; q0 += 32) {
                        const int q = q0 + lane;
                        const double2 a = __ldcs(src + 2 * q), b = __ldcs(src + 2 * q + 1);   // streamed: read once
                        const float4 v = make_float4((float)a.x, (float)a.y, (float)b.x, (float)b.y);   // :406-408
                        dst[q] = v;
                        chk(v);
                    }
                } else {
                    const float4* src = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(A.y) + (size_t)frame * n);
                    [[maybe_unused]] float4* dst = reinterpret_cast<float4*>(yscr + f * n);
#pragma unroll PROU
                    for (int q0 = 0; q0 < n / 4; q0 += 32) {
                        const int q = q0 + lane;
                        const float4 v = CONV ? __ldcs(src + q) : src[q];
                        if constexpr (CONV) dst[q] = v;
                        chk(v);
                    }
                }
                if (S.check_range && !__all_sync(FULL, ok) && lane == 0 && FPW * quad + f < A.B)
                    A.redo_list[atomicAdd(A.redo_count, 1)] = frame;
            }
            if constexpr (CONV) __syncwarp();          // the rows are written before anyone stages them
        }
#pragma unroll 1
        for (int blk = 0; blk < n / 8; ++blk) {
#if PCB_Q_PF64
            // CONV: the doubles of the next quad are on their way from DRAM to L2 during the last eighth of
            // this one (its prologue then waits for L2, not for DRAM; earlier they would only crowd L2)
            if (CONV && blk == n / 8 - n / 64 && next_quad < nquads) {
#pragma unroll 1
                for (int f = 0; f < FPW; f++) {
                    const int frame = min(FPW * next_quad + f, A.B - 1);
                    if (frame >= S.nf) {
                        const char* row = reinterpret_cast<const char*>(S.y64 + (size_t)(frame - S.nf) * n);
#pragma unroll 2
                        for (int l = lane; l < n / 16; l += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(row + 128 * l));
                    }
                }
            }
#endif
            // the 8 information flags of this subtree and, as bit 8, the first one of the next
            const uint32_t ib = __funnelshift_r(S.info_w[blk >> 2], S.info_w[(blk >> 2) + 1], 8 * (blk & 3)) & 0x1FFu;
            D.upper(blk, ybase);
            D.leaf8(blk, ib);
        }
        __syncwarp();
        // ---- the codeword estimates of every slot leave the kernel (setbit_last, d2 == M, put them in the LLR
        // region); the final selection (:432-473) is scl_select_kernel's (scl_select.cu): it runs once per frame, and
        // every SASS row of THIS kernel competes with the decoding loop for the instruction cache.
        constexpr int NW = G::NW, XHS = G::XHS;
#pragma unroll 1
        for (int fs = 0; fs < FPW; fs++) {
            const int frame = FPW * quad + fs;
            if (frame < A.B) {                          // (warp-uniform)
                const uint32_t* X = reinterpret_cast<const uint32_t*>(D.W + fs * G::FRAME_WORDS);
                uint32_t* dst = S.xhat + (size_t)frame * (LL * NW);
#pragma unroll 2
                for (int i0 = 0; i0 < LL * NW; i0 += 32) {
                    const int i = i0 + lane;
                    if (LL * NW >= 32 || i < LL * NW) dst[i] = X[(i / NW) * XHS + (i & (NW - 1))];
                }
                if constexpr (GEN) {
                    if (lane == 0 && ((redo_mask >> fs) & 1u)) atomicOr(&S.skip[frame >> 5], 1u << (frame & 31));
                }
            }
        }
        __syncwarp();
        quad = next_quad;
    }
}

template <int M, int LL, typename YT, int SRC>
static cudaError_t launch_mc(const Scl8qArgs& a, int blocks, cudaStream_t st) {
    const int W = q8::Geo<M, LL>::MW ? q8::Geo<M, LL>::WMAX : 1;      // `blocks` counts warps
    const size_t smem = (size_t)q8::Geo<M, LL>::TOTAL * W;
    cudaError_t e = cudaFuncSetAttribute(decode_scl8q_kernel<M, LL, YT, SRC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    decode_scl8q_kernel<M, LL, YT, SRC><<<(blocks + W - 1) / W, 32 * W, smem, st>>>(a);
    return cudaGetLastError();
}
template <int M, int LL, typename YT>
static cudaError_t launch_m(const Scl8qArgs& a, int blocks, cudaStream_t st) {
    if (a.gen) return launch_mc<M, LL, YT, 2>(a, blocks, st);
    return a.y64 ? launch_mc<M, LL, YT, 1>(a, blocks, st) : launch_mc<M, LL, YT, 0>(a, blocks, st);
}

template <int M, int LL, typename YT>
static cudaError_t occ_m(int* bps, int* warp_bytes) {
    const size_t smem = (size_t)q8::Geo<M, LL>::TOTAL;
    *warp_bytes = q8::Geo<M, LL>::TOTAL;
    const int W = q8::Geo<M, LL>::MW ? q8::Geo<M, LL>::WMAX : 1;
    cudaError_t e = cudaFuncSetAttribute(decode_scl8q_kernel<M, LL, YT, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem * W));
    if (e != cudaSuccess) return e;
    int nb = 0, nb2 = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, decode_scl8q_kernel<M, LL, YT, 1>, 32 * W, smem * W);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(decode_scl8q_kernel<M, LL, YT, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem * W));
    if (e != cudaSuccess) return e;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb2, decode_scl8q_kernel<M, LL, YT, 0>, 32 * W, smem * W);
    if (nb2 < nb) nb = nb2;                  // one grid size (and scratch size) for both input paths
    *bps = nb * W;                           // warps per SM
    return e;
}

// list sizes with a lane-per-path kernel: 32 / L frames per warp
bool scl8q_supported(int m, int L) { return m >= 8 && m <= 11 && (L == 2 || L == 4 || L == 8 || L == 16 || L == 32); }
int scl8q_frames_per_warp(int L) { return 32 / L; }

#ifdef PCB_Q_ONLY          // kernel experiments: one instantiation compiles in seconds
#define PCB_FOR_ALL(X) X(10, 8)
#else
#define PCB_FOR_ALL(X) \
    X(8, 8) X(9, 8) X(10, 8) X(11, 8) X(8, 16) X(9, 16) X(10, 16) X(11, 16) X(8, 32) X(9, 32) X(10, 32) X(11, 32) \
    X(8, 4) X(9, 4) X(10, 4) X(11, 4) X(8, 2) X(9, 2) X(10, 2) X(11, 2)
#endif

// info_w: host copy of the information-mask words (ceil(n/32)), passed as kernel parameters
// y64 != null: frames >= nf of the batch are doubles at y64 + (frame - nf) * n and yscr holds blocks * (32 / L) * n
// floats; check_range: fill the redo list (a.redo_count / a.redo_list) with the frames outside the safe input range
// gen != null: the frames are generated by the kernel itself (Scl8qGen, decode_launch.h) and counted, a.y is unused
cudaError_t launch_decode_scl8q(const DecodeArgs& a, int L, const uint32_t* info_w, uint32_t* xhat, void* l3_scratch, int* next_quad,
                                const double* y64, int nf, float* yscr, int check_range, const Scl8qGen* gen, int blocks, cudaStream_t st) {
    Scl8qArgs s;
    s.a = a;
    s.y64 = y64; s.nf = y64 ? nf : a.B; s.yscr = yscr; s.check_range = check_range;
    if ((y64 || gen) && !yscr) return cudaErrorInvalidValue;
    s.seed = 0; s.first_frame = 0; s.sigma = 0.f; s.gcrc = nullptr; s.gcrc_c0 = 0; s.skip = nullptr; s.mc_redo = nullptr; s.gen = gen ? 1 : 0;
    s.mc_redo_cap = 0; s.force_redo_mod = 0;
    if (gen) {
        if (!gen->skip || !gen->mc_redo || (a.crc && !gen->gcrc)) return cudaErrorInvalidValue;
        s.seed = gen->seed; s.first_frame = gen->first_frame; s.sigma = gen->sigma; s.gcrc = gen->gcrc; s.gcrc_c0 = gen->gcrc_c0;
        s.skip = gen->skip; s.mc_redo = gen->mc_redo; s.mc_redo_cap = gen->mc_redo_cap; s.force_redo_mod = gen->force_redo_mod;
        s.y64 = nullptr; s.check_range = 0;
    }
    s.next_quad = next_quad;
    s.l3 = reinterpret_cast<float4*>(l3_scratch);
    s.xhat = xhat;
    if (!xhat) return cudaErrorInvalidValue;
    for (int i = 0; i < 65; i++) s.info_w[i] = i < a.n / 32 ? info_w[i] : 0;
    if (!a.y_f32 && !y64 && !gen) return cudaErrorInvalidValue;   // a.y is float; doubles come in through y64
#define PCB_CASE(MM, LLL) if (a.m == MM && L == LLL) return launch_m<MM, LLL, float>(s, blocks, st);
    PCB_FOR_ALL(PCB_CASE)
#undef PCB_CASE
    return cudaErrorInvalidValue;
}

// bytes of global scratch one block needs (layer-3/4 arrays of its frames)
size_t scl8q_scratch_per_block(int m, int L) {
#define PCB_CASE(MM, LLL) if (m == MM && L == LLL) return (size_t)q8::Geo<MM, LLL>::SCRATCH_BLOCK_FLOAT4 * 16;
    PCB_FOR_ALL(PCB_CASE)
#undef PCB_CASE
    return 0;
}

// warps per block: all resident warps of an SM for the variants with several warps per block, else one
int scl8q_warps_per_block(int m, int L) {
#define PCB_CASE(MM, LLL) if (m == MM && L == LLL) return q8::Geo<MM, LLL>::MW ? q8::Geo<MM, LLL>::WMAX : 1;
    PCB_FOR_ALL(PCB_CASE)
#undef PCB_CASE
    return 1;
}

// resident warps per SM
cudaError_t scl8q_occupancy(int m, int L, int* bps, int* warp_bytes) {
#define PCB_CASE(MM, LLL) if (m == MM && L == LLL) return occ_m<MM, LLL, float>(bps, warp_bytes);
    PCB_FOR_ALL(PCB_CASE)
#undef PCB_CASE
    return cudaErrorInvalidValue;
}

}  // namespace pcb
