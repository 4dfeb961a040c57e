#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace pcb {

struct GenArgs {
    int n, m, k, crc, B;
    int with_shuffle;            // 1 for decoders 1,2 (PolarCodes.cpp:157 `decoder != 3`)
    int tx_is_u;                 // 1: tx_bits = u[n] (decoders 1,2), 0: xinf[k+crc] (decoder 3)
    int tx_words;
    float sigma;
    uint64_t seed, first_frame;
    const uint32_t* info_mask;
    const uint32_t* frozen_val;
    const uint16_t* info_pos;
    float* y;                    // [B][n]
    uint32_t* tx_bits;           // [B][tx_words]
};

cudaError_t launch_encode(const int* d_u, int* d_x, int n, int m, int with_shuffle, int B, int sm_count,
                          cudaStream_t st);
cudaError_t launch_generate(const GenArgs& g, int sm_count, cudaStream_t st);
// y64[count] -> y32[count], round to nearest (what SCLDecoderMinsum.cpp:406-408 does per frame)
cudaError_t launch_convert_f64_f32(const double* y64, float* y32, size_t count, int sm_count, cudaStream_t st);
cudaError_t launch_count(const uint32_t* dec, const uint32_t* tx, int words, int nbits, int B,
                         unsigned long long* counters, cudaStream_t st);

}  // namespace pcb
