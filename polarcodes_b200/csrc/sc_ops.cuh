// f / g of the two SC decoders and the register-resident size-S subtree, shared by the lane-per-frame
// kernels (decode_sc32.cu, decode_sct.cu).
#pragma once
#include "decode_common.cuh"

namespace pcb {
namespace s32 {

// f and g of the decoder: min-sum (SCDecoderMinsum.cpp:14-26) or exact (SCDecoder.cpp:19-36, double only)
template <bool EXACT, typename T>
__device__ __forceinline__ T opf(T a, T b) {
    if constexpr (EXACT) return (T)f_exact((double)a, (double)b);
    else return f_minsum<T>(a, b);
}
template <bool EXACT, typename T>
__device__ __forceinline__ T opg(T a, T b, int c) {
    if constexpr (EXACT) return (T)g_exact((double)a, (double)b, c);
    else return g_minsum<T>(a, b, c);
}

// The same f and g with the instruction count the hot loops of decode_sct.cu want (fp64 f: 7 instead of 10 SASS rows,
// three instead of six on the fp64 pipe; g: 3 instead of 5).  Same IEEE results: min of the magnitudes by the same
// compare, the sign from the same real product (an underflowing product counts as "+", SCDecoderMinsum.cpp:14-16),
// g = b + (+-a) with the sign flip done on the bit pattern (exact).  signbit = partial-sum bit at bit 31.
__device__ __forceinline__ double f_minsum_fast(double a, double b) {
    const int ah = __double2hiint(a), bh = __double2hiint(b), al = __double2loint(a), bl = __double2loint(b);
    const double p = a * b;
    const bool lt = fabs(b) < fabs(a);               // (the compare takes |.| as operand modifiers: no abs is materialised)
    int mh = (lt ? bh : ah) & 0x7fffffff;
    const int ml = lt ? bl : al;
    if (p < 0.) mh |= (int)0x80000000;
    return __hiloint2double(mh, ml);
}
__device__ __forceinline__ float f_minsum_fast(float a, float b) {
    const float p = a * b;
    uint32_t mn = __float_as_uint(fabsf(b) < fabsf(a) ? b : a) & 0x7fffffffu;
    if (p < 0.f) mn |= 0x80000000u;
    return __uint_as_float(mn);
}
__device__ __forceinline__ double g_sign(double a, double b, uint32_t signbit) {
    return b + __hiloint2double(__double2hiint(a) ^ (int)signbit, __double2loint(a));
}
__device__ __forceinline__ float g_sign(float a, float b, uint32_t signbit) { return b + __uint_as_float(__float_as_uint(a) ^ signbit); }
// 8-bit min-sum (decode_sc8.cu; the reference's sketch SCDecoderMinsum.cpp:17-30, g saturating): values in [-127, 127]
__device__ __forceinline__ int f_minsum_fast(int a, int b) {
    const int mm = min(abs(a), abs(b));
    return ((a ^ b) < 0) ? -mm : mm;
}
__device__ __forceinline__ int g_sign(int a, int b, uint32_t signbit) {
    const int v = b + (signbit ? -a : a);
    return max(-127, min(127, v));
}
// one copy of the tanh / atanh sequence per kernel (inlined at every call site the exact-SC tile kernel is 1 MB of SASS)
#ifndef PCB_EXACT_INLINE
static __device__ __noinline__ double f_exact_call(double a, double b) { return f_exact(a, b); }
#else
__device__ __forceinline__ double f_exact_call(double a, double b) { return f_exact(a, b); }
#endif
template <bool EXACT, typename T>
__device__ __forceinline__ T opf2(T a, T b) {
    if constexpr (EXACT) return (T)f_exact_call((double)a, (double)b);
    else return f_minsum_fast(a, b);
}
// SCDecoder.cpp:27-36 is b == 0 ? y1 + y2 : y2 - y1: the same sum with the sign of y1 flipped
template <bool EXACT, typename T>
__device__ __forceinline__ T opg2(T a, T b, uint32_t signbit) { return g_sign(a, b, signbit); }

// SC decoding of a subtree of S leaves held in registers (SCDecoderMinsum.cpp:32-66): returns its
// partial sums, the decisions in phase order in ub.  ib/fb: information flags / frozen values.
template <int S, bool EXACT, typename T>
__device__ __forceinline__ uint32_t sc_node(const T (&a)[S], uint32_t ib, uint32_t fb, uint32_t& ub) {
    if constexpr (S == 1) {
        const uint32_t u = (ib & 1u) ? (a[0] > (T)0 ? 0u : 1u) : (fb & 1u);   // :33-47, an LLR of exactly 0 decodes to 1
        ub = u;
        return u;
    } else {
        constexpr int Hf = S / 2;
        T l[Hf], r[Hf];
#pragma unroll
        for (int j = 0; j < Hf; j++) l[j] = opf2<EXACT, T>(a[j], a[j + Hf]);
        uint32_t uL, uR;
        const uint32_t pL = sc_node<Hf, EXACT, T>(l, ib, fb, uL);
#pragma unroll
        for (int j = 0; j < Hf; j++) r[j] = opg2<EXACT, T>(a[j], a[j + Hf], (pL << (31 - j)) & 0x80000000u);
        const uint32_t pR = sc_node<Hf, EXACT, T>(r, ib >> Hf, fb >> Hf, uR);
        ub = uL | (uR << Hf);
        return (pL ^ pR) | (pR << Hf);
    }
}

// smallest magnitude for which no product of two LLRs underflows (the all-information shortcut below needs it)
template <typename T>
struct Tiny;
template <>
struct Tiny<double> {
    static __device__ __forceinline__ bool ok(double a) { return fabs(a) >= 0x1p-511; }    // false for 0 and NaN
};
template <>
struct Tiny<float> {
    static __device__ __forceinline__ bool ok(float a) { return fabsf(a) >= 0x1p-63f; }
};

// The same subtree with the two prunings of decode_sct.cu applied inside it (the pattern is warp-uniform, so the
// tests are uniform branches): an all-frozen (zeros) subtree costs nothing, an all-information subtree is the hard
// decisions of its inputs -- exact unless an input is zero or tiny, which sets `bad` (the frame is decoded again
// by the generic kernel).
template <>
struct Tiny<int> {
    static __device__ __forceinline__ bool ok(int a) { return a != 0; }
};
template <int S, bool EXACT, typename T, bool R1 = true>
__device__ __forceinline__ uint32_t sc_node_p(const T (&a)[S], uint32_t ib, uint32_t fb, uint32_t& ub, bool& bad) {
    if constexpr (S == 1) {
        const uint32_t u = (ib & 1u) ? (a[0] > (T)0 ? 0u : 1u) : (fb & 1u);
        ub = u;
        return u;
    } else {
        constexpr uint32_t ALL = (1u << S) - 1;
        if (((ib | fb) & ALL) == 0u) { ub = 0u; return 0u; }
        if (R1 && (ib & ALL) == ALL) {
            uint32_t beta = 0;
            bool ok = true;
#pragma unroll
            for (int j = 0; j < S; j++) {
                ok = ok && Tiny<T>::ok(a[j]);
                beta |= (a[j] > (T)0 ? 0u : 1u) << j;
            }
            bad = bad || !ok;
            uint32_t u = beta;                       // u = beta F^(x)s, packed
            u ^= (u >> 1) & (0x55u & ALL);
            if constexpr (S >= 4) u ^= (u >> 2) & (0x33u & ALL);
            if constexpr (S >= 8) u ^= (u >> 4) & 0x0Fu;
            ub = u;
            return beta;
        }
        constexpr int Hf = S / 2;
        constexpr uint32_t HALF = (1u << Hf) - 1;
        T l[Hf], r[Hf];
        uint32_t uL = 0, uR, pL = 0;
        if (((ib | fb) & HALF) != 0u) {
#pragma unroll
            for (int j = 0; j < Hf; j++) l[j] = opf2<EXACT, T>(a[j], a[j + Hf]);
            pL = sc_node_p<Hf, EXACT, T, R1>(l, ib, fb, uL, bad);
        }
#pragma unroll
        for (int j = 0; j < Hf; j++) r[j] = opg2<EXACT, T>(a[j], a[j + Hf], (pL << (31 - j)) & 0x80000000u);
        const uint32_t pR = sc_node_p<Hf, EXACT, T, R1>(r, ib >> Hf, fb >> Hf, uR, bad);
        ub = uL | (uR << Hf);
        return (pL ^ pR) | (pR << Hf);
    }
}

}  // namespace s32
}  // namespace pcb
