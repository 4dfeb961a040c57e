// debug: the packed 8-bit f / g of decode_sc8.cu and the int instantiation of sc_node_p against scalar code, on the device
#include <cstdio>
#include <cstdint>
#include "../../polarcodes_b200/csrc/decode_common.cuh"
#include "../../polarcodes_b200/csrc/sc_ops.cuh"
namespace pcb { namespace sc8 {
// prmt with the full selector nibbles (bit 3 = replicate the sign of the selected byte; __byte_perm keeps only 3 bits)
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
__device__ __forceinline__ uint32_t smask(uint32_t a) { return prmt(a, 0u, 0xba98u); }
__device__ __forceinline__ uint32_t negm(uint32_t x, uint32_t m) { const uint32_t t = x ^ m; return (((t & 0x7f7f7f7fu) + (m & 0x01010101u)) ^ (t & 0x80808080u)); }
__device__ __forceinline__ uint32_t f4(uint32_t a, uint32_t b) {
    const uint32_t sa = smask(a), sb = smask(b);
    const uint32_t aa = (a ^ sa) + (sa & 0x01010101u), ab = (b ^ sb) + (sb & 0x01010101u);
    const uint32_t ge = smask((aa | 0x80808080u) - ab);
    const uint32_t mn = (ab & ge) | (aa & ~ge);
    return negm(mn, sa ^ sb);
}
__device__ __forceinline__ uint32_t g4(uint32_t a, uint32_t b, uint32_t nib) {
    const uint32_t m = ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;
    const uint32_t an = negm(a, m);
    const uint32_t alo = prmt(an, 0u, 0x9180u), ahi = prmt(an, 0u, 0xb3a2u);
    const uint32_t blo = prmt(b, 0u, 0x9180u), bhi = prmt(b, 0u, 0xb3a2u);
    const uint32_t slo = __vmaxs2(__vmins2(__vadd2(alo, blo), 0x007f007fu), 0xff81ff81u);
    const uint32_t shi = __vmaxs2(__vmins2(__vadd2(ahi, bhi), 0x007f007fu), 0xff81ff81u);
    return prmt(slo, shi, 0x6420u);
}
}}
__device__ int f8(int a, int b) { int mm = min(abs(a), abs(b)); return ((a ^ b) & 0x80) == 0 ? mm : -mm; }
__device__ int g8(int a, int b, int c) { int v = b + a * (1 - 2 * c); return max(-127, min(127, v)); }
__global__ void k(int* bad) {
    int a = (int)blockIdx.x - 127, b = (int)threadIdx.x - 127;
    if (b > 127) return;
    int va[4] = {a, b, -a, 0}, vb[4] = {b, a, b, a};
    uint32_t A = 0, B = 0;
    for (int j = 0; j < 4; j++) { A |= (uint32_t)(va[j] & 0xFF) << (8 * j); B |= (uint32_t)(vb[j] & 0xFF) << (8 * j); }
    uint32_t r = pcb::sc8::f4(A, B);
    for (int j = 0; j < 4; j++) if ((int)(signed char)(r >> (8 * j)) != f8(va[j], vb[j])) atomicAdd(bad, 1);
    for (uint32_t nib = 0; nib < 16; nib++) {
        r = pcb::sc8::g4(A, B, nib);
        for (int j = 0; j < 4; j++) if ((int)(signed char)(r >> (8 * j)) != g8(va[j], vb[j], (nib >> j) & 1)) atomicAdd(bad + 1, 1);
    }
    if (pcb::s32::opf2<false, int>(a, b) != f8(a, b)) atomicAdd(bad + 2, 1);
    if (pcb::s32::opg2<false, int>(a, b, 0x80000000u) != g8(a, b, 1) || pcb::s32::opg2<false, int>(a, b, 0u) != g8(a, b, 0)) atomicAdd(bad + 3, 1);
}
int main() {
    int* d; cudaMalloc(&d, 16); cudaMemset(d, 0, 16);
    k<<<255, 256>>>(d);
    int h[4]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("mismatches: f4 %d g4 %d opf2<int> %d opg2<int> %d  (%s)\n", h[0], h[1], h[2], h[3], cudaGetErrorString(cudaGetLastError()));
    return 0;
}
