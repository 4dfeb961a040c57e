// Tuned list-8 SCL min-sum decoder (SCLDecoderMinsum.cpp:225-478 semantics, bit-exact with
// decode_generic.cu): ONE WARP DECODES ONE FRAME, lane = sub*8 + p with p the path slot (0..7)
// and sub (0..3) the quarter of every LLR layer that lane owns.
//
// Layout choices:
//  * LLR element j of a layer belongs to sub = j mod 4.  The half-split pair (j, j+s) is then
//    lane-local on every layer of size >= 4, so a whole f/g chain from a large layer down to
//    size 4 needs no shuffle and no __syncwarp; the chunk of a lane is contiguous in shared
//    memory ("sub-major" order) and moves as float4 / float2.
//  * layers 0 (channel), 1 and 2 are never stored: layer 3 is computed straight from y with
//    the three top levels fused (every n/8 phases).  y is converted once per event and staged
//    in the shared-memory region of layers 4.. (dead at those phases).  This cuts the LLR
//    storage of a frame from 32 KB to 9.5 KB, i.e. 16 resident frames per SM instead of 6.
//  * layers of size 4, 2, 1 live in registers; a clone copies them with shuffles.
//  * lazy copy: slot p writes its own arrays; a per-path word holds, per layer, the slot to
//    read from (3 bits per layer).  A clone copies that word.
//  * prune: ~85 % of the calls exit early (every penalised candidate above every unpenalised
//    one); otherwise 16 candidates x 16 = 256 compares = 8 per lane on the metric bit patterns.
#include "decode_common.cuh"
#include "decode_launch.h"

namespace pcb {

namespace s8 {

constexpr unsigned FULL = 0xffffffffu;

// Tables in constant memory, built at compile time (one warp per block: per-block shared-memory
// copies would cost 1.5 KB of every frame's slice).
struct LutTable { unsigned char v[1024]; };
struct CrcTable { unsigned short v[256]; };
// lut[mask*4 + j] = position of the j-th HIGHEST set bit of the 8-bit mask
constexpr LutTable make_lut() {
    LutTable t{};
    for (int i = 0; i < 1024; i++) {
        unsigned bm = (unsigned)i >> 2;
        for (int j = 0; j < (i & 3); j++) {
            if (bm) { int hb = 7; while (!((bm >> hb) & 1)) hb--; bm &= ~(1u << hb); }
        }
        int hb = 0;
        if (bm) { hb = 7; while (!((bm >> hb) & 1)) hb--; }
        t.v[i] = (unsigned char)hb;
    }
    return t;
}
// Resources.cpp:29-62 (poly 0x31) and :73-106 (poly 0x1021) regenerated from the polynomials
constexpr CrcTable make_crc(int bits) {
    CrcTable t{};
    for (int i = 0; i < 256; i++) {
        unsigned r = bits == 8 ? (unsigned)i : ((unsigned)i << 8);
        for (int b = 0; b < 8; b++) {
            if (bits == 8) r = (r & 0x80) ? ((r << 1) ^ 0x31) & 0xFF : (r << 1) & 0xFF;
            else r = (r & 0x8000) ? ((r << 1) ^ 0x1021) & 0xFFFF : (r << 1) & 0xFFFF;
        }
        t.v[i] = (unsigned short)r;
    }
    return t;
}
__constant__ LutTable c_lut = make_lut();
__constant__ CrcTable c_crc8 = make_crc(8);
__constant__ CrcTable c_crc16 = make_crc(16);

template <int M>
struct Geo {
    static constexpr int N = 1 << M;
    static constexpr int LTOP = 3;          // first stored LLR layer (size N/8)
    static constexpr int LBOT = M - 3;      // last stored LLR layer (size 8); sizes 4,2,1 are registers
    static constexpr int NW = N / 32;       // partial-sum words per slot (heap: layer of size S at bits [S,2S))
    static constexpr int HS = NW + 1;       // padded slot stride of the heap (bank spread)
    __host__ __device__ static constexpr int c(int lam) { return (N >> lam) / 4; }
    __host__ __device__ static constexpr int cs(int lam) { return (c(lam) % 8 == 0) ? c(lam) + 4 : c(lam); }
    __host__ __device__ static constexpr int off(int lam) {
        int o = 0;
        for (int l = LTOP; l < lam; ++l) o += 32 * cs(l);
        return o;
    }
    static constexpr int PFLOATS = off(LBOT + 1);
    static constexpr bool STAGE = (PFLOATS - off(LTOP + 1)) >= N;   // y fits in the region of layers 4..
    static constexpr int XW = 8 * HS;                      // final codeword estimates (alias the LLR region)
    static constexpr int EW_MAX = N / 32;
    static constexpr int OFF_P = 0;
    static constexpr int P_BYTES = ((PFLOATS > XW + 8 * EW_MAX ? PFLOATS : XW + 8 * EW_MAX) * 4 + 15) & ~15;
    static constexpr int OFF_H = P_BYTES;
    static constexpr int OFF_KEYS = OFF_H + ((8 * HS * 4 + 15) & ~15);
    static constexpr int TOTAL = OFF_KEYS + 64;
};

// f for frames whose channel values are all 0 or within [2^-40, 2^100] (checked per frame, see
// the kernel): every LLR is then an integer multiple of 2^-63 below 2^111, so no sum overflows,
// no NaN appears and no product of two non-zero LLRs underflows.  Under these conditions
// min(|a|,|b|) * (a*b < 0 ? -1 : 1) (SCLDecoderMinsum.cpp:40-42) equals fminf(|a|,|b|) with the
// sign bit of the product (a zero result may come out as -0, which nothing can observe).
__device__ __forceinline__ float ff(float a, float b) {
    const float m = fminf(fabsf(a), fabsf(b));
    const float pr = a * b;
    return __uint_as_float(__float_as_uint(m) | (__float_as_uint(pr) & 0x80000000u));
}
// g with the partial-sum bit already moved to the sign position (0 or 0x80000000)
__device__ __forceinline__ float gs(float a, float b, uint32_t signbit) {
    return b + __uint_as_float(__float_as_uint(a) ^ signbit);
}
__device__ __forceinline__ uint32_t sgn(uint32_t w, int pos) { return (w << (31 - pos)) & 0x80000000u; }
__device__ __forceinline__ float4 ff4(float4 a, float4 b) {
    return make_float4(ff(a.x, b.x), ff(a.y, b.y), ff(a.z, b.z), ff(a.w, b.w));
}
// bits of the 4 elements at positions pos, pos+4, pos+8, pos+12 of cw
__device__ __forceinline__ float4 gs4(float4 a, float4 b, uint32_t cw, int pos) {
    return make_float4(gs(a.x, b.x, sgn(cw, pos)), gs(a.y, b.y, sgn(cw, pos + 4)), gs(a.z, b.z, sgn(cw, pos + 8)),
                       gs(a.w, b.w, sgn(cw, pos + 12)));
}

// Per-path state that a fork can replace (a clone takes its parent's).
struct PathState {
    float met;
    uint32_t pidx, cidx, w0;
    float A2, A1_0, A1_1;
    int u;
};

// Everything of continuePaths_UnfrozenBit (SCLDecoderMinsum.cpp:315-400) except the fast exit:
// growth while the list is not full, else the median select; then the clones take over their
// parents' state.  One copy in the binary (arguments and result travel in registers).
__device__ __noinline__ PathState fork_slow(PathState s, float A0, int lane, int ninfo, uint32_t* keys) {
    const int sub = lane >> 3, p = lane & 7;
    const float pf0 = s.met + fmaxf(-A0, 0.f);   // pen(llr,0) = |llr| iff llr < 0   (:308-313)
    const float pf1 = s.met + fmaxf(A0, 0.f);    // pen(llr,1) = |llr| iff llr > 0
    uint32_t r;   // u | take<<1 | srcp<<2
    if (ninfo < 3) {
        // growth (:329-341): live slots are [8-nact, 8); parent q clones into slot 2*pbase-1-q
        const int nact = 1 << ninfo, pbase = 8 - nact;
        const bool take = p >= pbase - nact && p < pbase;
        r = take ? (3u | ((uint32_t)(2 * pbase - 1 - p) << 2)) : ((uint32_t)p << 2);
    } else {
        // keys are stored complemented: key < kc  <=>  carry-out of (kc + ~key)
        if (sub == 0) *reinterpret_cast<uint2*>(keys + 2 * p) = make_uint2(~__float_as_uint(pf0), ~__float_as_uint(pf1));
        __syncwarp();
        const uint32_t kc = __float_as_uint((sub & 1) ? pf1 : pf0);
        const uint4 ka = *reinterpret_cast<const uint4*>(keys + 8 * (sub >> 1));
        const uint4 kb = *reinterpret_cast<const uint4*>(keys + 8 * (sub >> 1) + 4);
        int cnt = 0;   // number of keys below kc, accumulated with carry chains (IADD3 + IADD3.X)
#define PCB_LT(x) asm("{ .reg .u32 t; add.cc.u32 t, %1, %2; addc.u32 %0, %0, 0; }" : "+r"(cnt) : "r"(x), "r"(kc));
        PCB_LT(ka.x) PCB_LT(ka.y) PCB_LT(ka.z) PCB_LT(ka.w) PCB_LT(kb.x) PCB_LT(kb.y) PCB_LT(kb.z) PCB_LT(kb.w)
#undef PCB_LT
        cnt += __shfl_xor_sync(FULL, cnt, 16);
        unsigned bal = __ballot_sync(FULL, cnt < 8) & 0xFFFFu;
        if (__popc(bal) != 8) {                // a tie straddles the median (:362-371)
            const bool in = (bal >> (lane & 15)) & 1;
            const unsigned tkey = __reduce_max_sync(FULL, in ? kc : 0u);
            const unsigned tb = __ballot_sync(FULL, in && kc == tkey) & 0xFFFFu;
            const int need = 8 - (__popc(bal) - __popc(tb));
            const unsigned t0 = tb & 0xFF, t1 = tb >> 8, lt = (1u << p) - 1;
            const int before = __popc(t0 & lt) + __popc(t1 & lt) + ((sub & 1) ? ((t0 >> p) & 1) : 0);
            bal = __ballot_sync(FULL, in && (kc < tkey || before < need)) & 0xFFFFu;
        }
        __syncwarp();
        const unsigned k0m = bal & 0xFF, k1m = bal >> 8;
        const unsigned nonem = ~(k0m | k1m) & 0xFF, bothm = k0m & k1m;
        // killed slots are pushed ascending (:372-376) and popped by the cloning parents in ascending
        // order (:382-389): the j-th killed slot receives the clone of the j-th HIGHEST parent
        const uint32_t src = c_lut.v[bothm * 4 + (__popc(nonem & ((1u << p) - 1)) & 3)];
        r = ((nonem >> p) & 1) ? (3u | (src << 2)) : ((((k0m >> p) & 1) ? 0u : 1u) | ((uint32_t)p << 2));
    }
    s.u = r & 1;
    s.met = s.u ? pf1 : pf0;
    // clones copy the parent's state (:145-156 without copying arrays)
    __syncwarp();   // orders the parent's shared-memory writes before the clone's lazy reads
    const int sl = (lane & 24) | ((r >> 2) & 7);
    const float m1 = __shfl_sync(FULL, pf1, sl);
    const uint32_t pi = __shfl_sync(FULL, s.pidx, sl);
    const uint32_t ci = __shfl_sync(FULL, s.cidx, sl);
    const uint32_t ww = __shfl_sync(FULL, s.w0, sl);
    const float t2 = __shfl_sync(FULL, s.A2, sl);
    const float t0 = __shfl_sync(FULL, s.A1_0, sl), t1 = __shfl_sync(FULL, s.A1_1, sl);
    if (r & 2u) { s.met = m1; s.pidx = pi; s.cidx = ci; s.w0 = ww; s.A2 = t2; s.A1_0 = t0; s.A1_1 = t1; }
    return s;
}

template <int M, typename YT>
struct Dec {
    using G = Geo<M>;
    int lane, sub, p;
    float* P;          // warp slice: LLR layers
    uint32_t* H;       // partial-sum heaps
    uint32_t* keys;
    const YT* y;
    float met;
    uint32_t pidx, cidx, w0;
    float A2, A1_0, A1_1, A0;
    int ninfo;         // information bits seen so far (warp-uniform): the list is full from the 4th on

    __device__ __forceinline__ float* chunk(int lam, int slot) const {
        return P + G::off(lam) + (sub * 8 + slot) * G::cs(lam);
    }
    __device__ __forceinline__ int pfield(int lam) const { return (pidx >> (3 * (lam - G::LTOP))) & 7; }
    __device__ __forceinline__ int cfield(int lam) const { return (cidx >> (3 * (lam - 1))) & 7; }
    // word of partial-sum layer `lam` (size >= 32) holding this lane's elements k in [8*w8, 8*w8+8)
    // of the sub-array starting at element `base`: element k's bit ends up at position 4*(k&7)
    __device__ __forceinline__ uint32_t cword(int lam, int base, int w8) const {
        return H[cfield(lam) * G::HS + ((G::N >> lam) >> 5) + (base >> 5) + w8] >> sub;
    }

    // ---- layer 3 straight from the channel values (layers 0, 1, 2 are virtual) ---------
    // ev = i0 / (N/8) in 0..7: bit 2 -> layer 1 is a g-layer, bit 1 -> layer 2, bit 0 -> layer 3.
    // L1[x] = op1(y[x], y[x+4E]), L2[x] = op2(L1[x], L1[x+2E]), L3[j] = op3(L2[j], L2[j+E]), E = N/8.
    __device__ __forceinline__ void top3(int ev) {
        constexpr int E = G::N / 8, C3 = G::c(3), E4 = E / 4;
        const bool g1 = ev & 4, g2 = ev & 2, g3 = ev & 1;
        float* Y = P + G::off(G::LTOP + 1);      // staging area: [sub][N/4] floats, element j at [j&3][j>>2]
        if (G::STAGE) {
            __syncwarp();                        // everyone is done with layers 4.. of the previous subtree
            for (int q = lane; q < G::N / 4; q += 32) {
                float v0, v1, v2, v3;
                if (sizeof(YT) == 8) {
                    const double2 a = *reinterpret_cast<const double2*>(y + 4 * q);
                    const double2 b = *reinterpret_cast<const double2*>(y + 4 * q + 2);
                    v0 = (float)a.x; v1 = (float)a.y; v2 = (float)b.x; v3 = (float)b.y;   // :406-408
                } else {
                    const float4 a = *reinterpret_cast<const float4*>(y + 4 * q);
                    v0 = a.x; v1 = a.y; v2 = a.z; v3 = a.w;
                }
                Y[q] = v0; Y[G::N / 4 + q] = v1; Y[2 * (G::N / 4) + q] = v2; Y[3 * (G::N / 4) + q] = v3;
            }
            __syncwarp();
        }
        const float* ys = Y + sub * (G::N / 4);
        float* dst = chunk(3, p);
#pragma unroll 1
        for (int k0 = 0; k0 < C3; k0 += 8) {
            uint32_t c1[4] = {0, 0, 0, 0}, c2[2] = {0, 0}, c3w = 0;
            if (g1) {
#pragma unroll
                for (int t = 0; t < 4; t++) c1[t] = cword(1, t * E, k0 >> 3);
            }
            if (g2) { c2[0] = cword(2, 0, k0 >> 3); c2[1] = cword(2, E, k0 >> 3); }
            if (g3) c3w = cword(3, 0, k0 >> 3);
#pragma unroll
            for (int kk = 0; kk < 8; kk += 4) {
                float4 yv[8];
#pragma unroll
                for (int t = 0; t < 8; t++) {
                    if (G::STAGE) yv[t] = *reinterpret_cast<const float4*>(ys + t * E4 + k0 + kk);
                    else {
                        const int j = sub + 4 * (k0 + kk) + t * E;
                        yv[t] = make_float4((float)y[j], (float)y[j + 4], (float)y[j + 8], (float)y[j + 12]);
                    }
                }
                float4 l1[4], l2[2], o;
                if (g1) {
#pragma unroll
                    for (int t = 0; t < 4; t++) l1[t] = gs4(yv[t], yv[t + 4], c1[t], 4 * kk);
                } else {
#pragma unroll
                    for (int t = 0; t < 4; t++) l1[t] = ff4(yv[t], yv[t + 4]);
                }
                if (g2) { l2[0] = gs4(l1[0], l1[2], c2[0], 4 * kk); l2[1] = gs4(l1[1], l1[3], c2[1], 4 * kk); }
                else { l2[0] = ff4(l1[0], l1[2]); l2[1] = ff4(l1[1], l1[3]); }
                o = g3 ? gs4(l2[0], l2[1], c3w, 4 * kk) : ff4(l2[0], l2[1]);
                *reinterpret_cast<float4*>(dst + k0 + kk) = o;
            }
        }
        if (G::STAGE) __syncwarp();              // staging area is overwritten by layer 4 next
    }

    // ---- one stored layer from the previous stored layer ------------------------------
    // ISG: g-layer (first layer of a chain, source slot is the lazily shared one)
    template <int LAM, bool ISG>
    __device__ __forceinline__ void layer() {
        constexpr int C = G::c(LAM);          // outputs per lane
        const int slot = ISG ? pfield(LAM - 1) : p;
        const float* src = chunk(LAM - 1, slot);
        float* dst = chunk(LAM, p);
        if constexpr (C >= 8) {
#pragma unroll 1
            for (int k0 = 0; k0 < C; k0 += 8) {
                uint32_t cw = 0;
                if (ISG) cw = cword(LAM, 0, k0 >> 3);
#pragma unroll
                for (int kk = 0; kk < 8; kk += 4) {
                    const float4 a = *reinterpret_cast<const float4*>(src + k0 + kk);
                    const float4 b = *reinterpret_cast<const float4*>(src + C + k0 + kk);
                    *reinterpret_cast<float4*>(dst + k0 + kk) = ISG ? gs4(a, b, cw, 4 * kk) : ff4(a, b);
                }
            }
        } else if constexpr (C == 4) {        // size 16: bits 16 + sub + 4k of w0
            const float4 a = *reinterpret_cast<const float4*>(src);
            const float4 b = *reinterpret_cast<const float4*>(src + 4);
            *reinterpret_cast<float4*>(dst) = ISG ? gs4(a, b, w0 >> (16 + sub), 0) : ff4(a, b);
        } else {                              // C == 2, size 8: bits 8 + sub + 4k of w0
            const float4 ab = *reinterpret_cast<const float4*>(src);   // a0 a1 b0 b1
            float2 o;
            if (ISG) {
                const uint32_t cw = w0 >> (8 + sub);
                o.x = gs(ab.x, ab.z, sgn(cw, 0)); o.y = gs(ab.y, ab.w, sgn(cw, 4));
            } else { o.x = ff(ab.x, ab.z); o.y = ff(ab.y, ab.w); }
            *reinterpret_cast<float2*>(dst) = o;
        }
    }

    // g at layer ll (one of LTOP+1..LBOT); every g variant appears once in the binary
    template <int LAM>
    __device__ __forceinline__ void g_dispatch(int ll) {
        if constexpr (LAM > G::LTOP) {
            if (ll == LAM) layer<LAM, true>();
            else g_dispatch<LAM - 1>(ll);
        }
    }
    // f layers ll+1..LBOT; every f variant appears once in the binary
    template <int LAM>
    __device__ __forceinline__ void f_below(int ll) {
        if constexpr (LAM <= G::LBOT) {
            if (ll < LAM) layer<LAM, false>();
            f_below<LAM + 1>(ll);
        }
    }

    // LLR work at the first phase i0 = 8*blk of a size-8 subtree, down to layer LBOT
    __device__ __forceinline__ void upper(int blk) {
        const int z = blk ? (__ffs(blk) - 1) : 31;
        int ll;
        if (z >= M - 6) { top3(blk >> (M - 6)); ll = G::LTOP; }     // i0 is a multiple of N/8
        else { ll = G::LBOT - z; g_dispatch<G::LBOT>(ll); }
        f_below<G::LTOP + 1>(ll);
        // layers >= ll are now this slot's own
        const uint32_t low = (1u << (3 * (ll - G::LTOP))) - 1;
        pidx = (pidx & low) | ((0x9249249u * (uint32_t)p) & ~low);
    }

    // ---- register layers ------------------------------------------------------------
    __device__ __forceinline__ void a1_from_a2(bool isg) {
        const float x = __shfl_xor_sync(FULL, A2, 16);
        float own;
        if (isg) {
            const float lo = (sub < 2) ? A2 : x, hi = (sub < 2) ? x : A2;
            own = gs(lo, hi, sgn(w0, 2 + (sub & 1)));
        } else own = ff(A2, x);
        const float oth = __shfl_xor_sync(FULL, own, 8);
        A1_0 = (sub & 1) ? oth : own;
        A1_1 = (sub & 1) ? own : oth;
    }
    __device__ __forceinline__ void a2_from_a3(bool isg) {
        const int slot = isg ? pfield(G::LBOT) : p;
        const float2 v = *reinterpret_cast<const float2*>(chunk(G::LBOT, slot));
        A2 = isg ? gs(v.x, v.y, sgn(w0, 4 + sub)) : ff(v.x, v.y);
    }

    // ---- fork / prune of one information bit (SCLDecoderMinsum.cpp:315-400) ----------
    __device__ __forceinline__ int fork() {
        if (ninfo >= 3) {
            // Fast exit (no kill, no clone): the candidates of path p are {met, met + |llr|}.  If
            // every penalised candidate is strictly above every unpenalised one, the L smallest
            // are exactly the L unpenalised ones (no tie possible), each path keeps its likely
            // bit and its metric.  ~85 % of the prune calls at the operating point end here.
            const float hi = met + fmaxf(fabsf(A0), 0.f);          // NaN llr -> penalty 0 (as :308-313)
            const unsigned minhi = __reduce_min_sync(FULL, __float_as_uint(hi));
            if (__all_sync(FULL, __float_as_uint(met) < minhi)) return A0 < 0.f ? 1 : 0;   // max met < min hi
        }
        PathState s{met, pidx, cidx, w0, A2, A1_0, A1_1, 0};
        s = fork_slow(s, A0, lane, ninfo, keys);
        met = s.met; pidx = s.pidx; cidx = s.cidx; w0 = s.w0; A2 = s.A2; A1_0 = s.A1_0; A1_1 = s.A1_1;
        if (ninfo < 3) ninfo++;
        return s.u;
    }

    // decision at phase i0+T with the leaf LLR in A0
    template <int T>
    __device__ __forceinline__ int decide(uint32_t infobyte, uint32_t fvbyte) {
        if ((infobyte >> T) & 1) return fork();
        (void)fvbyte;                   // the tuned kernel is only used when every frozen value is 0
        met += fmaxf(-A0, 0.f);         // :413-418
        return 0;
    }

    // partial-sum combine at the last phase of a subtree (t = 7): d2 = 3 + ctz(blk+1)
    __device__ __forceinline__ void setbit_last(int blk, int u) {
        const int d2 = min(M, 3 + __ffs(blk + 1) - 1);
        uint32_t T = (uint32_t)u;
        T = (((w0 >> 1) ^ T) & 1u) | (T << 1);
        T = (((w0 >> 2) ^ T) & 3u) | (T << 2);
        T = (((w0 >> 4) ^ T) & 15u) | (T << 4);
        if (d2 == 3) { w0 = (w0 & ~0xFF00u) | (T << 8); return; }
        T = (((w0 >> 8) ^ T) & 0xFFu) | (T << 8);
        if (d2 == 4) { w0 = (w0 & 0x0000FFFFu) | (T << 16); return; }
        T = (((w0 >> 16) ^ T) & 0xFFFFu) | (T << 16);
        const int D = 1 << d2;
        // destination layer m-d2 sits at heap words [D/32, 2D/32); layer 0 (the final codeword
        // estimate, written at the very last phase) goes to the then-dead LLR region instead
        uint32_t* hp = (d2 == M) ? (reinterpret_cast<uint32_t*>(P) + p * G::HS - G::NW) : (H + p * G::HS);
        __syncwarp();
        if (sub == 0) hp[2 * D / 32 - 1] = T;
        __syncwarp();
        for (int w = 32; w < D; w <<= 1) {
            const int lgw = 31 - __clz(w), layc = M - lgw, nw = w >> 5;
            const uint32_t* cs = H + cfield(layc) * G::HS + nw;
            for (int kk = sub; kk < nw; kk += 4)
                hp[(2 * D - 2 * w) / 32 + kk] = cs[kk] ^ hp[(2 * D - w) / 32 + kk];
            __syncwarp();
        }
        if (d2 < M) {
            const int sh = 3 * (M - d2 - 1);
            cidx = (cidx & ~(7u << sh)) | ((uint32_t)p << sh);
        }
    }

    // the 8 phases of one size-8 subtree; layer LBOT of this subtree is already computed
    __device__ __forceinline__ void leaf8(int blk, uint32_t ib, uint32_t fb) {
        int u;
        // t = 0
        a2_from_a3(false); a1_from_a2(false); A0 = ff(A1_0, A1_1);
        u = decide<0>(ib, fb); w0 = (w0 & ~2u) | ((uint32_t)u << 1);
        // t = 1
        A0 = gs(A1_0, A1_1, sgn(w0, 1));
        u = decide<1>(ib, fb);
        { uint32_t T = (((w0 >> 1) ^ u) & 1u) | ((uint32_t)u << 1); w0 = (w0 & ~0xCu) | (T << 2); }
        // t = 2
        a1_from_a2(true); A0 = ff(A1_0, A1_1);
        u = decide<2>(ib, fb); w0 = (w0 & ~2u) | ((uint32_t)u << 1);
        // t = 3
        A0 = gs(A1_0, A1_1, sgn(w0, 1));
        u = decide<3>(ib, fb);
        { uint32_t T = (((w0 >> 1) ^ u) & 1u) | ((uint32_t)u << 1); T = (((w0 >> 2) ^ T) & 3u) | (T << 2);
          w0 = (w0 & ~0xF0u) | (T << 4); }
        // t = 4
        a2_from_a3(true); a1_from_a2(false); A0 = ff(A1_0, A1_1);
        u = decide<4>(ib, fb); w0 = (w0 & ~2u) | ((uint32_t)u << 1);
        // t = 5
        A0 = gs(A1_0, A1_1, sgn(w0, 1));
        u = decide<5>(ib, fb);
        { uint32_t T = (((w0 >> 1) ^ u) & 1u) | ((uint32_t)u << 1); w0 = (w0 & ~0xCu) | (T << 2); }
        // t = 6
        a1_from_a2(true); A0 = ff(A1_0, A1_1);
        u = decide<6>(ib, fb); w0 = (w0 & ~2u) | ((uint32_t)u << 1);
        // t = 7
        A0 = gs(A1_0, A1_1, sgn(w0, 1));
        u = decide<7>(ib, fb);
        setbit_last(blk, u);
    }
};

}  // namespace s8

// Kernel parameters: the generic arguments plus the frozen pattern itself.  Reading the mask
// words from the constant bank with a loop-counter index keeps every branch on them provably
// warp-uniform (no divergence bookkeeping around the shuffles).
struct Scl8Args {
    DecodeArgs a;
    uint32_t info_w[64];   // n <= 2048
    uint32_t fv_w[64];
};

// One warp per block (blockDim.x == 32): the frame index depends on blockIdx only.
template <int M, typename YT>
__global__ void __launch_bounds__(32, 20) decode_scl8_kernel(const __grid_constant__ Scl8Args S) {
    const DecodeArgs& A = S.a;
    using G = s8::Geo<M>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x;
    unsigned char* ws = smem_raw;
    s8::Dec<M, YT> D;
    D.lane = lane; D.sub = lane >> 3; D.p = lane & 7;
    D.P = reinterpret_cast<float*>(ws + G::OFF_P);
    D.H = reinterpret_cast<uint32_t*>(ws + G::OFF_H);
    D.keys = reinterpret_cast<uint32_t*>(ws + G::OFF_KEYS);
    const int n = G::N;
    const int EW = (A.stride + 31) >> 5;

    for (int frame = blockIdx.x; frame < A.B; frame += gridDim.x) {
        D.y = reinterpret_cast<const YT*>(A.y) + (size_t)frame * n;
        {   // range check that licenses the arithmetic shortcuts of ff(): |y| == 0 or in [2^-40, 2^100]
            bool ok = true;
#pragma unroll 1
            for (int q = lane; q < n / 4; q += 32) {
                float v[4];
                if (sizeof(YT) == 8) {
                    const double2 a = *reinterpret_cast<const double2*>(D.y + 4 * q);
                    const double2 b = *reinterpret_cast<const double2*>(D.y + 4 * q + 2);
                    v[0] = (float)a.x; v[1] = (float)a.y; v[2] = (float)b.x; v[3] = (float)b.y;
                } else {
                    const float4 a = *reinterpret_cast<const float4*>(D.y + 4 * q);
                    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
                }
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float a = fabsf(v[e]);
                    ok = ok && (a == 0.f || (a >= 0x1p-40f && a <= 0x1p100f));
                }
            }
            if (!__all_sync(FULL, ok)) {   // hand the frame to the generic (reference-literal) kernel
                if (lane == 0) A.redo_list[atomicAdd(A.redo_count, 1)] = frame;
                continue;
            }
        }
        D.met = 0.f; D.pidx = 0; D.cidx = 0; D.w0 = 0; D.ninfo = 0;
        D.A2 = D.A1_0 = D.A1_1 = D.A0 = 0.f;
        uint32_t iw = 0, fw = 0;
#pragma unroll 1
        for (int blk = 0; blk < n / 8; ++blk) {
            iw = S.info_w[blk >> 2]; fw = S.fv_w[blk >> 2];
            D.upper(blk);
            D.leaf8(blk, (iw >> (8 * (blk & 3))) & 0xFF, (fw >> (8 * (blk & 3))) & 0xFF);
        }
        __syncwarp();
        // ---- final selection (:432-473): u-hat = x-hat F^(x)m per slot, CRC, first pass wins ----
        uint32_t* X = reinterpret_cast<uint32_t*>(D.P);          // x-hat of every slot (setbit_last, d2 == M)
        uint32_t* E = X + G::XW;
        constexpr int NW = G::NW, HS = G::HS, lgN = M - 5;
        const int q0 = A.crc ? 0 : 7;   // without a CRC the reference returns slot L-1 unconditionally
#pragma unroll 1
        for (int e = lane + (q0 << lgN); e < (8 << lgN); e += 32) {
            const int q = e >> lgN, a = e & (NW - 1);
            uint32_t x = X[q * HS + a];
            x ^= (x >> 1) & 0x55555555u; x ^= (x >> 2) & 0x33333333u; x ^= (x >> 4) & 0x0F0F0F0Fu;
            x ^= (x >> 8) & 0x00FF00FFu; x ^= (x >> 16) & 0x0000FFFFu;
            X[q * HS + a] = x;
        }
        __syncwarp();
#pragma unroll 1
        for (int sw = 1; sw < NW; sw <<= 1) {
#pragma unroll 1
            for (int e = lane + (q0 << lgN); e < (8 << lgN); e += 32) {
                const int q = e >> lgN, a = e & (NW - 1);
                if (!(a & sw)) X[q * HS + a] ^= X[q * HS + a + sw];
            }
            __syncwarp();
        }
#pragma unroll 1
        for (int q = q0; q < 8; q++) {
#pragma unroll 1
            for (int t0 = 0; t0 < A.stride; t0 += 32) {
                const int t = t0 + lane;
                int bit = 0;
                if (t < A.stride) { const int pos = A.info_pos[t]; bit = (X[q * HS + (pos >> 5)] >> (pos & 31)) & 1; }
                const unsigned wd = __ballot_sync(FULL, bit);
                if (lane == 0) E[q * EW + (t0 >> 5)] = wd;
            }
        }
        __syncwarp();
        bool ok = false;
        if (A.crc && lane < 8) {
            const uint32_t* e = E + lane * EW;
            unsigned r = A.crc == 8 ? 0xFFu : 0xFFFFu;
#pragma unroll 1
            for (int b = 0; b < A.k / 8; b++) {
                const unsigned byte = __brev((e[b >> 2] >> (8 * (b & 3))) & 0xFFu) >> 24;
                if (A.crc == 8) r = s8::c_crc8.v[(r ^ byte) & 0xFF];
                else r = ((r << 8) ^ s8::c_crc16.v[((r >> 8) ^ byte) & 0xFF]) & 0xFFFF;
            }
            unsigned rx = 0;
            for (int t = 0; t < A.crc; t++) { const int q = A.k + t; rx = (rx << 1) | ((e[q >> 5] >> (q & 31)) & 1); }
            ok = (rx == r);
        }
        const unsigned okm = __ballot_sync(FULL, ok);
        const int chosen = okm ? (__ffs(okm) - 1) : 7;
        if (A.sol)
            for (int t = lane; t < A.stride; t += 32)
                A.sol[(size_t)frame * A.stride + t] = (E[chosen * EW + (t >> 5)] >> (t & 31)) & 1;
        if (A.sol_bits)
            for (int t = lane; t < EW; t += 32) A.sol_bits[(size_t)frame * EW + t] = E[chosen * EW + t];
        __syncwarp();
    }
}

template <int M, typename YT>
static cudaError_t launch_m(const Scl8Args& a, int blocks, cudaStream_t st) {
    const size_t smem = (size_t)s8::Geo<M>::TOTAL;
    cudaError_t e = cudaFuncSetAttribute(decode_scl8_kernel<M, YT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    decode_scl8_kernel<M, YT><<<blocks, 32, smem, st>>>(a);
    return cudaGetLastError();
}

template <int M, typename YT>
static cudaError_t occ_m(int* bps, int* warp_bytes) {
    const size_t smem = (size_t)s8::Geo<M>::TOTAL;
    *warp_bytes = s8::Geo<M>::TOTAL;
    cudaError_t e = cudaFuncSetAttribute(decode_scl8_kernel<M, YT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, decode_scl8_kernel<M, YT>, 32, smem);
}

bool scl8_supported(int m) { return m >= 8 && m <= 11; }

// info_w / fv_w: host copies of the mask words (ceil(n/32) each), passed as kernel parameters
cudaError_t launch_decode_scl8(const DecodeArgs& a, const uint32_t* info_w, const uint32_t* fv_w, int blocks,
                               cudaStream_t st) {
    Scl8Args s;
    s.a = a;
    for (int i = 0; i < 64; i++) { s.info_w[i] = i < a.n / 32 ? info_w[i] : 0; s.fv_w[i] = i < a.n / 32 ? fv_w[i] : 0; }
#define PCB_CASE(MM) \
    case MM: return a.y_f32 ? launch_m<MM, float>(s, blocks, st) : launch_m<MM, double>(s, blocks, st);
    switch (a.m) { PCB_CASE(8) PCB_CASE(9) PCB_CASE(10) PCB_CASE(11) }
#undef PCB_CASE
    return cudaErrorInvalidValue;
}

// resident blocks (= warps = frames) per SM of the more register-hungry (double input) variant
cudaError_t scl8_occupancy(int m, int* bps, int* warp_bytes) {
    switch (m) {
        case 8: return occ_m<8, double>(bps, warp_bytes);
        case 9: return occ_m<9, double>(bps, warp_bytes);
        case 10: return occ_m<10, double>(bps, warp_bytes);
        case 11: return occ_m<11, double>(bps, warp_bytes);
    }
    return cudaErrorInvalidValue;
}

}  // namespace pcb
