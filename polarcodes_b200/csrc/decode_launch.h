// Host/device-shared layout of one warp's shared-memory slice and the kernel launchers.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "decode_common.cuh"

namespace pcb {

struct WarpLayout {
    int off_P, off_H, off_pidx, off_cidx, off_keys, off_E, total;
};

// LLR layers first_smem_layer..m (L slots each), the partial-sum heaps (2n bits per slot),
// the published per-slot index words, the 2L candidate keys and the packed solution words.
__host__ __device__ inline WarpLayout warp_layout(int sizeofT, int L, int n, int ls, int stride, int /*mode*/) {
    WarpLayout w;
    auto al = [](int x) { return (x + 15) & ~15; };
    const int HW = (2 * n >= 32) ? (2 * n / 32) : 1;
    const int EW = (stride + 31) >> 5;
    int o = 0;
    w.off_P = o;    o = al(o + sizeofT * L * ((n >> (ls - 1)) - 1));
    w.off_H = o;    o = al(o + 4 * L * HW);
    w.off_pidx = o; o = al(o + 8 * L);
    w.off_cidx = o; o = al(o + 8 * L);
    w.off_keys = o; o = al(o + 4 * 2 * L);
    w.off_E = o;    o = al(o + 4 * L * EW);
    w.total = o;
    return w;
}

// bytes of global scratch one warp needs for LLR layers 1..ls-1
inline size_t warp_scratch_bytes(int sizeofT, int L, int n, int ls) {
    return (size_t)sizeofT * L * (size_t)(n - (n >> (ls - 1)));
}

// Monte-Carlo mode of the list kernel (decode_scl8q.cu): the kernel generates, decodes and counts
struct Scl8qGen {
    uint64_t seed, first_frame;        // Philox key, global id of frame 0 of the launch
    float sigma;
    const uint32_t* gcrc;              // device: [ceil((k+crc)/32)][16] words of the payload -> CRC map (crc > 0)
    uint32_t gcrc_c0;                  // CRC of the all-zero payload
    uint32_t* skip;                    // device: ceil(B/32) zeroed words; the kernel sets the bit of every frame it puts on mc_redo
    unsigned long long* mc_redo;       // device: [0] count, [1..] global ids of the frames the kernel could not decode itself
    int mc_redo_cap;
    int force_redo_mod;                // test hook
};

// Final selection of the lane-per-path list kernel (scl_select.cu)
struct SelectArgs {
    const uint32_t* xhat;              // [B][L][n/32] codeword estimates of every slot (written by decode_scl8q_kernel)
    int B, L, n, k, crc, stride;
    const uint32_t* info_mask;         // n/32 words, device
    const uint32_t* synd;              // crc > 0: 16 rows of n/32 words, see pcb_set_code
    uint32_t synd_c0;
    int* sol;                          // [B][stride] one int per bit, or null
    uint32_t* sol_bits;                // [B][ceil(stride/32)] packed, or null
    // Monte-Carlo mode: count the errors of the k information bits against the regenerated payload
    int count;
    const uint32_t* skip;              // frames with their bit set are not counted (they are decoded again elsewhere)
    unsigned long long* counters;      // [0] += frame errors, [1] += bit errors
    uint64_t seed, first_frame;
};
cudaError_t launch_scl_select(const SelectArgs& a, int sm_count, cudaStream_t st);

cudaError_t launch_decode_generic(const DecodeArgs& a, int L, int mode, int fp64, int blocks, int wpb,
                                  cudaStream_t st);

// pcb_decode_batch_dev_noalloc: size of the file-scope scratch array of decode_generic.cu and its device
// address on the current device (cudaGetSymbolAddress, cached per device)
#ifndef PCB_NOALLOC_SCRATCH_BYTES
#define PCB_NOALLOC_SCRATCH_BYTES (8u << 20)      // = include/polar_b200.h
#endif
cudaError_t noalloc_scratch_address(unsigned char** ptr, size_t* bytes);

}  // namespace pcb
