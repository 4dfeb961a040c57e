// Frame generation shared by generate_kernel (aux_kernels.cu) and the list kernel's generating prologue
// (decode_scl8q.cu): the random stream of a frame is a function of (seed, global frame id, index) only, and
// both places run the very same device functions, so a frame is the same bits and the same floats
// wherever it is made (tests compare pcb_generate_batch + decode with pcb_montecarlo counter by counter).
//   payload bits      PolarCodes.cpp:133-135   Philox word w of the frame, tag 'info'
//   Transform / Awgn  Encoder.cpp:13-15,35-41  y = (c == 0 ? 1 : -1) + sigma * z, z from Philox + Box-Muller
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace pcb {

// ---------------------------------------------------------------- Philox4x32-10 ---------
// PCB_PHILOX_COMPACT (decode_scl8q.cu: code size is instruction-cache traffic there): one out-of-line copy with
// a rolled round loop.  Integer arithmetic: the same words either way.
#ifdef PCB_PHILOX_COMPACT
#define PCB_PHILOX_ATTR __noinline__
constexpr int kPhiloxUnroll = 1;
#else
#define PCB_PHILOX_ATTR __forceinline__
constexpr int kPhiloxUnroll = 10;
#endif
struct Philox {
    uint32_t k0, k1;
    __device__ PCB_PHILOX_ATTR uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
        uint32_t a = k0, b = k1;
#pragma unroll kPhiloxUnroll
        for (int r = 0; r < 10; r++) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            const uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            a += 0x9E3779B9u; b += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};

// Box-Muller on the SFU: lg2 / sqrt / sin / cos approximations (relative error ~1e-6, far below the
// 24-bit grid of u2).  The channel of the Monte-Carlo mode is statistical by construction (the reference
// draws from mt19937 + std::normal_distribution); its moments and tails are tested on the GPU.
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
    // u1 in (0, 1] on the full 32-bit grid near zero ((float)a is exact for a < 2^24, which is where the
    // tail of the normal lives): the smallest value 2^-33 gives |z| up to 6.76; u2 on a 24-bit grid
    const float u1 = a ? __uint2float_rn(a) * (1.0f / 4294967296.0f) : (0.5f / 4294967296.0f);
    const float ang = ((float)(b >> 8) + 0.5f) * (6.283185307179586f / 16777216.0f);   // in (0, 2 pi)
    float r;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(-1.3862943611198906f * __log2f(u1)));   // sqrt(-2 ln u1)
    z0 = r * __cosf(ang); z1 = r * __sinf(ang);
}

// payload word w (bits 32w .. 32w+31 of the k information bits; bits at or above k are zero)
__device__ __forceinline__ uint32_t payload_word(const Philox& ph, uint32_t f_lo, uint32_t f_hi, int w, int k) {
    const uint4 r = ph(f_lo, f_hi, (uint32_t)(w >> 2), 0x696E666Fu);
    uint32_t v = (w & 3) == 0 ? r.x : (w & 3) == 1 ? r.y : (w & 3) == 2 ? r.z : r.w;
    const int hi = k - w * 32;
    if (hi < 32) v = hi > 0 ? (v & ((1u << hi) - 1)) : 0u;
    return v;
}

// channel values 4q .. 4q+3 of the frame; bits4: the four codeword bits (bit t = element 4q + t)
__device__ __forceinline__ float4 channel4(const Philox& ph, uint32_t f_lo, uint32_t f_hi, int q, uint32_t bits4, float sigma) {
    const uint4 r = ph(f_lo, f_hi, (uint32_t)q, 0x6E6F6973u);
    float z[4];
    box_muller(r.x, r.y, z[0], z[1]);
    box_muller(r.z, r.w, z[2], z[3]);
    float4 v;
    v.x = ((bits4 & 1u) ? -1.f : 1.f) + sigma * z[0];
    v.y = ((bits4 & 2u) ? -1.f : 1.f) + sigma * z[1];
    v.z = ((bits4 & 4u) ? -1.f : 1.f) + sigma * z[2];
    v.w = ((bits4 & 8u) ? -1.f : 1.f) + sigma * z[3];
    return v;
}

}  // namespace pcb
