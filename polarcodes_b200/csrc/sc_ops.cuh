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

// The same f and g with the instruction count the hot loops of decode_sct.cu want (fp64 f: 8 instead of 10 SASS rows,
// three instead of six on the fp64 pipe; g: 3 instead of 5).  Same IEEE results: min of the magnitudes by the same
// compare, the sign from the same real product (an underflowing product counts as "+", SCDecoderMinsum.cpp:14-16),
// g = b + (+-a) with the sign flip done on the bit pattern (exact).  signbit = partial-sum bit at bit 31.
__device__ __forceinline__ double f_minsum_fast(double a, double b) {
    const int ah = __double2hiint(a), bh = __double2hiint(b), al = __double2loint(a), bl = __double2loint(b);
    const double p = a * b;
    const int aah = ah & 0x7fffffff, abh = bh & 0x7fffffff;
    const bool lt = __hiloint2double(abh, bl) < __hiloint2double(aah, al);
    int mh = lt ? abh : aah;
    const int ml = lt ? bl : al;
    if (p < 0.) mh |= (int)0x80000000;
    return __hiloint2double(mh, ml);
}
__device__ __forceinline__ float f_minsum_fast(float a, float b) {
    const float p = a * b;
    const uint32_t aa = __float_as_uint(a) & 0x7fffffffu, ab = __float_as_uint(b) & 0x7fffffffu;
    uint32_t mn = (__uint_as_float(ab) < __uint_as_float(aa)) ? ab : aa;
    if (p < 0.f) mn |= 0x80000000u;
    return __uint_as_float(mn);
}
__device__ __forceinline__ double g_sign(double a, double b, uint32_t signbit) {
    return b + __hiloint2double(__double2hiint(a) ^ (int)signbit, __double2loint(a));
}
__device__ __forceinline__ float g_sign(float a, float b, uint32_t signbit) { return b + __uint_as_float(__float_as_uint(a) ^ signbit); }
template <bool EXACT, typename T>
__device__ __forceinline__ T opf2(T a, T b) {
    if constexpr (EXACT) return (T)f_exact((double)a, (double)b);
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

}  // namespace s32
}  // namespace pcb
