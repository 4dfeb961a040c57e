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

cudaError_t launch_decode_generic(const DecodeArgs& a, int L, int mode, int fp64, int blocks, int wpb,
                                  cudaStream_t st);

}  // namespace pcb
