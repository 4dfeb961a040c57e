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
    int* redo_count;             // not null: range check for the tuned list kernel (what scl8_prepass_kernel does), frames
    int* redo_list;              // outside the safe input range are appended here
};

cudaError_t launch_encode(const int* d_u, int* d_x, int n, int m, int with_shuffle, int B, int sm_count,
                          cudaStream_t st);
cudaError_t launch_encode_packed(const uint32_t* d_u, uint32_t* d_x, int n, int m, int with_shuffle, int B, int sm_count, cudaStream_t st);
cudaError_t launch_generate(const GenArgs& g, int sm_count, cudaStream_t st);
// pre-pass of the list-8 quad kernel: y64 (may be null) -> y32 conversion and the range check that
// fills the redo list (see aux_kernels.cu)
cudaError_t launch_scl8_prepass(const double* y64, float* y32, int n, int B, int nf, int* redo_count, int* redo_list,
                                int sm_count, cudaStream_t st);
cudaError_t launch_count(const uint32_t* dec, const uint32_t* tx, int words, int nbits, int B,
                         unsigned long long* counters, cudaStream_t st);

}  // namespace pcb
