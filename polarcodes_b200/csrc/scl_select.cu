// Final selection of the list decoder (SCLDecoderMinsum.cpp:432-473) for the lane-per-path kernel, as its own
// kernel: decode_scl8q_kernel leaves the codeword estimate x-hat of every slot in global memory ([B][L][n/32]
// words) and this kernel, one warp per frame, picks the slot, computes u-hat = x-hat F^(x)m for that slot only and
// writes the information (+CRC) bits.  It used to be the tail of the decode kernel; it runs once per frame, its 330
// SASS rows competed with the decoding loop of the 15 other warps of an SM for the instruction cache (the decode
// kernel sits on a code-size cliff), and moving 1 KB per frame through L2 is free next to that.
//
//  * slot order, no metric sort, no active check: the FIRST slot whose CRC passes wins; slot L-1 if none passes
//    or there is no CRC (:466-472).
//  * CRC check as a syndrome of the codeword estimate: crc(info bits of x F) ^ (appended CRC bits of x F) is
//    GF(2)-affine in x, so slot q passes iff S x_q == c0 with S a crc x n bit matrix that depends on the code only
//    (row t as n/32 words at synd[t * n/32 ..), built by pcb_set_code) and c0 the value for the all-zero word:
//    16 AND/POPC parities per word and one warp XOR reduction per slot instead of transform + compress + CRC per slot.
//  * lane l owns words l (and l + 32) of a codeword estimate: the transform inside a word is five shift-xor steps,
//    across words five shuffles; the information bits of a word are compressed (Hacker's Delight 7-4) and OR-ed into
//    the packed output at the bit offset that the information counts of the words before it give.
//  * Monte-Carlo mode (count != 0): the frame-error compare of the loop (PolarCodes.cpp:181-186) against the payload,
//    regenerated from (seed, global frame id) with the generator's own functions (gen_common.cuh); bit errors too.
#include "decode_launch.h"
#define PCB_PHILOX_COMPACT 1
#include "gen_common.cuh"

namespace pcb {

template <int WPL>
__global__ void __launch_bounds__(128) scl_select_kernel(const SelectArgs A) {
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ uint32_t sel_smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int NW = A.n >> 5, EW = (A.stride + 31) >> 5;
    uint32_t* Ec = sel_smem + wib * (EW + 1);
    uint32_t sm[WPL][16];                // this lane's words of the syndrome rows (zero rows above crc)
    uint32_t mk[WPL];
    int off[WPL];
    {
        int base = 0;
#pragma unroll
        for (int w = 0; w < WPL; w++) {
            const bool in = lane + 32 * w < NW;
#pragma unroll
            for (int t = 0; t < 16; t++) sm[w][t] = (A.crc > t && in) ? __ldg(A.synd + t * NW + lane + 32 * w) : 0u;
            mk[w] = in ? __ldg(A.info_mask + lane + 32 * w) : 0u;
            const int c = __popc(mk[w]);
            int inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += v; }
            off[w] = base + inc - c;
            base += __shfl_sync(FULL, inc, 31);
        }
    }
    const Philox ph{(uint32_t)A.seed, (uint32_t)(A.seed >> 32)};
    unsigned long long fe_acc = 0, be_acc = 0;
    for (int frame = blockIdx.x * wpb + wib; frame < A.B; frame += gridDim.x * wpb) {
        const uint32_t* X = A.xhat + (size_t)frame * ((size_t)A.L * NW);
        int chosen = A.L - 1;
        if (A.crc) {
#pragma unroll 1
            for (int q = 0; q < A.L; q++) {
                uint32_t acc = 0;
#pragma unroll
                for (int w = 0; w < WPL; w++) {
                    const uint32_t x = lane + 32 * w < NW ? X[q * NW + lane + 32 * w] : 0u;
#pragma unroll
                    for (int t = 0; t < 16; t++) acc ^= (uint32_t)(__popc(x & sm[w][t]) & 1) << t;
                }
                if (__reduce_xor_sync(FULL, acc) == A.synd_c0) { chosen = q; break; }
            }
        }
        // u-hat of the chosen slot: transform, then the information bits compressed into Ec
        uint32_t xs[WPL];
#pragma unroll
        for (int w = 0; w < WPL; w++) {
            uint32_t x = lane + 32 * w < NW ? X[chosen * NW + lane + 32 * w] : 0u;
            x ^= (x >> 1) & 0x55555555u; x ^= (x >> 2) & 0x33333333u; x ^= (x >> 4) & 0x0F0F0F0Fu;
            x ^= (x >> 8) & 0x00FF00FFu; x ^= (x >> 16) & 0x0000FFFFu;
#pragma unroll
            for (int sw = 1; sw < 32; sw <<= 1) {
                const uint32_t y = __shfl_xor_sync(FULL, x, sw);
                if (sw < NW && !(lane & sw)) x ^= y;
            }
            xs[w] = x;
        }
        if constexpr (WPL == 2) xs[0] ^= xs[1];
        for (int t = lane; t <= EW; t += 32) Ec[t] = 0;
        __syncwarp();
#pragma unroll
        for (int w = 0; w < WPL; w++) {
            uint32_t m = mk[w], x = xs[w] & m, mkk = ~m << 1;
#pragma unroll
            for (int i = 0; i < 5; i++) {
                uint32_t mp = mkk ^ (mkk << 1);
                mp ^= mp << 2; mp ^= mp << 4; mp ^= mp << 8; mp ^= mp << 16;
                const uint32_t mv = mp & m;
                m = (m ^ mv) | (mv >> (1 << i));
                const uint32_t t = x & mv;
                x = (x ^ t) | (t >> (1 << i));
                mkk &= ~mp;
            }
            const int c = __popc(mk[w]), o = off[w];
            if (c) {
                atomicOr(&Ec[o >> 5], x << (o & 31));
                if ((o & 31) + c > 32) atomicOr(&Ec[(o >> 5) + 1], x >> (32 - (o & 31)));
            }
        }
        __syncwarp();
        if (A.sol)
            for (int t = lane; t < A.stride; t += 32) A.sol[(size_t)frame * A.stride + t] = (Ec[t >> 5] >> (t & 31)) & 1;
        if (A.sol_bits)
            for (int t = lane; t < EW; t += 32) A.sol_bits[(size_t)frame * EW + t] = Ec[t];
        if (A.count && !((A.skip[frame >> 5] >> (frame & 31)) & 1u)) {
            const uint64_t fid = A.first_frame + (uint64_t)frame;
            int be = 0;
            for (int w = lane; w * 32 < A.k; w += 32) {
                uint32_t x = Ec[w] ^ payload_word(ph, (uint32_t)fid, (uint32_t)(fid >> 32), w, A.k);
                const int hi = A.k - 32 * w;
                if (hi < 32) x &= (1u << hi) - 1;                      // the CRC bits behind the payload are not counted
                be += __popc(x);
            }
            be = __reduce_add_sync(FULL, be);
            fe_acc += be > 0; be_acc += (unsigned)be;
        }
        __syncwarp();
    }
    if (A.count && lane == 0 && (fe_acc | be_acc)) { atomicAdd(&A.counters[0], fe_acc); atomicAdd(&A.counters[1], be_acc); }
}

cudaError_t launch_scl_select(const SelectArgs& a, int sm_count, cudaStream_t st) {
    if (!a.xhat || !a.info_mask || (a.crc && !a.synd) || a.n < 32 || a.n > 2048) return cudaErrorInvalidValue;
    if (a.count && (!a.skip || !a.counters)) return cudaErrorInvalidValue;
    if (a.B <= 0) return cudaSuccess;
    const int wpb = 4, EW = (a.stride + 31) >> 5;
    int blocks = (a.B + wpb - 1) / wpb;
    if (blocks > sm_count * 16) blocks = sm_count * 16;
    const size_t smem = (size_t)wpb * (EW + 1) * 4;
    if (a.n > 1024) scl_select_kernel<2><<<blocks, wpb * 32, smem, st>>>(a);
    else scl_select_kernel<1><<<blocks, wpb * 32, smem, st>>>(a);
    return cudaGetLastError();
}

}  // namespace pcb
