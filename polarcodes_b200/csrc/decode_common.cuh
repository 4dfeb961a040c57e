// Shared definitions for the batched SC / SCL decode kernels (sm_100a).
//
// Reference semantics (paths relative to /root/reference/PolarCodes/):
//   min-sum f/g            SCDecoderMinsum.cpp:14-16,24-26 (fp64), SCLDecoderMinsum.cpp:40-42,51-53 (fp32)
//   exact f/g              SCDecoder.cpp:15-36
//   path metric penalty    SCLDecoderMinsum.cpp:308-313
// All arithmetic here is plain IEEE (no fast-math, no FTZ, compiled with -fmad=false) so that
// every LLR is produced by the same rounded operations in the same order as on the CPU.
#pragma once
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

namespace pcb {

enum : int { MODE_SC_EXACT = 1, MODE_SC_MINSUM = 2, MODE_SCL = 3 };

struct DecodeArgs {
    int n, m, k, crc, B;
    int first_smem_layer;       // LLR layers >= this live in shared memory, lower ones in `scratch`
    int warp_smem_bytes;        // shared-memory slice per warp (one warp decodes one frame)
    int stride;                 // solution bits per frame: n (SC) or k+crc (SCL)
    const void* y;              // [B][n] channel values
    int y_f32;                  // 1: float, 0: double
    const double* y64_tail;     // not null (with y_f32 = 1): frame f >= y64_first is doubles at y64_tail + (f - y64_first) * n
    int y64_first;
    const uint32_t* info_mask;  // ceil(n/32) words, bit i set = information position
    const uint32_t* frozen_val; // ceil(n/32) words, frozen value at frozen positions
    const uint16_t* info_pos;   // [k+crc] ascending information positions (SCL)
    int* sol;                   // [B][stride] one int per bit, or nullptr
    uint32_t* sol_bits;         // [B][ceil(stride/32)] packed, or nullptr
    unsigned char* scratch;     // global scratch for LLR layers < first_smem_layer
    size_t scratch_per_warp;
    double sigma;               // exact SC: LLR = 2y/sigma^2
    // Out-of-range hand-over between the tuned list-8 kernel and the generic kernel: the tuned
    // kernel uses arithmetic shortcuts that are bit-exact only when every |y| is 0 or within
    // [2^-40, 2^100]; it appends any other frame to `redo_list` (count in *redo_count) and skips
    // it, and a follow-up launch of the generic kernel decodes exactly those frames.
    int* redo_count;            // device counter (tuned kernel: written; generic kernel: read)
    int* redo_list;             // device list of frame indexes
    int use_redo_list;          // generic kernel: decode frames redo_list[0 .. *redo_count) only
};

__device__ __forceinline__ float pcb_abs(float a) { return fabsf(a); }
__device__ __forceinline__ double pcb_abs(double a) { return fabs(a); }

// f(a,b) = min(|a|,|b|) * (a*b < 0 ? -1 : 1): the sign comes from a real product, so an
// underflowing product counts as "+" (SCDecoderMinsum.cpp:14-16).  std::min(x,y) = (y<x)?y:x.
template <typename T>
__device__ __forceinline__ T f_minsum(T a, T b) {
    T aa = pcb_abs(a), ab = pcb_abs(b);
    T mn = (ab < aa) ? ab : aa;
    T p = a * b;
    return (p < (T)0) ? -mn : mn;
}
// g(a,b,c) = b + a*(1-2c): the product is exact, one rounded add (SCDecoderMinsum.cpp:24-26).
template <typename T>
__device__ __forceinline__ T g_minsum(T a, T b, int c) {
    return c ? (b - a) : (b + a);
}
// SCDecoder.cpp:19-25 (libm tanh/atanh on the CPU; CUDA's differ in ulps: statistical parity)
__device__ __forceinline__ double f_exact(double y1, double y2) {
    double a = 2. * atanh(tanh(y1 / 2) * tanh(y2 / 2));
    if (isinf(a) || isnan(a)) {
        double aa = fabs(y1), ab = fabs(y2);
        double mn = (ab < aa) ? ab : aa;
        return (a < 0) ? -mn : mn;
    }
    return a;
}
// SCDecoder.cpp:27-36
__device__ __forceinline__ double g_exact(double y1, double y2, int b) { return b == 0 ? y1 + y2 : y2 - y1; }

// SCLDecoderMinsum.cpp:308-313
__device__ __forceinline__ float pm_penalty(float llr, int u) {
    return ((u == 0 && llr < 0.f) || (u == 1 && llr > 0.f)) ? fabsf(llr) : 0.f;
}

template <int L>
struct Log2 {
    static constexpr int value = 1 + Log2<L / 2>::value;
};
template <>
struct Log2<1> {
    static constexpr int value = 0;
};

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned r;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(r));
    return r;
}

}  // namespace pcb
