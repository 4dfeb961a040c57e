// Latency / throughput microbenchmarks of the warp primitives the list decoder leans on (sm_100a).
// usage: ./lat   (prints cycles per dependent op for one warp, and per op with 4 independent chains)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define REP 256

template <int MODE>
__global__ void k(uint32_t* out, int ilp) {
    __shared__ uint32_t sm[64 * 4];
    const int lane = threadIdx.x & 31;
    uint32_t x0 = lane * 7 + 1, x1 = lane * 5 + 3, x2 = lane * 3 + 5, x3 = lane + 9;
    const unsigned gmask = 0xFFu << (lane & 24);
    uint32_t* my = sm + (threadIdx.x >> 5) * 64;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < REP; it++) {
        if (MODE == 0) {          // 3-level xor butterfly min (8-lane groups)
#define BF(x) x = min(x, __shfl_xor_sync(0xffffffffu, x, 1)) + 1; x = min(x, __shfl_xor_sync(0xffffffffu, x, 2)) + 1; x = min(x, __shfl_xor_sync(0xffffffffu, x, 4)) + 1;
            BF(x0) if (ilp > 1) { BF(x1) } if (ilp > 2) { BF(x2) BF(x3) }
        } else if (MODE == 1) {   // full-mask redux (CREDUX), 1 per step
            x0 = __reduce_min_sync(0xffffffffu, x0) + lane;
            if (ilp > 1) x1 = __reduce_min_sync(0xffffffffu, x1) + lane;
            if (ilp > 2) { x2 = __reduce_min_sync(0xffffffffu, x2) + lane; x3 = __reduce_min_sync(0xffffffffu, x3) + lane; }
        } else if (MODE == 2) {   // 4 full-mask redux, one per group, then select
            uint32_t r = 0;
#pragma unroll
            for (int g = 0; g < 4; g++) {
                uint32_t v = __reduce_min_sync(0xffffffffu, ((lane >> 3) == g) ? x0 : 0xffffffffu);
                if ((lane >> 3) == g) r = v;
            }
            x0 = r + lane;
        } else if (MODE == 3) {   // smem: STS, syncwarp, 2x LDS.128, min tree
            my[lane] = x0;
            __syncwarp();
            const uint4 a = *reinterpret_cast<const uint4*>(my + (lane & 24));
            const uint4 b = *reinterpret_cast<const uint4*>(my + (lane & 24) + 4);
            __syncwarp();
            x0 = min(min(min(a.x, a.y), min(a.z, a.w)), min(min(b.x, b.y), min(b.z, b.w))) + lane;
        } else if (MODE == 4) {   // group-mask redux (what nvcc makes of it)
            x0 = __reduce_min_sync(gmask, x0) + lane;
        } else if (MODE == 5) {   // vote.all + dependent
            x0 += __all_sync(0xffffffffu, x0 > 5) ? 1 : 2;
        } else if (MODE == 6) {   // single shfl idx dependent
            x0 = __shfl_sync(0xffffffffu, x0, (x0 + 1) & 31) + 1;
            if (ilp > 1) x1 = __shfl_sync(0xffffffffu, x1, (x1 + 1) & 31) + 1;
            if (ilp > 2) { x2 = __shfl_sync(0xffffffffu, x2, (x2 + 1) & 31) + 1; x3 = __shfl_sync(0xffffffffu, x3, (x3 + 1) & 31) + 1; }
        } else if (MODE == 7) {   // ballot dependent
            x0 += __ballot_sync(0xffffffffu, x0 & 1);
        } else if (MODE == 8) {   // match_any on group id... (not used)
            x0 += __popc(__match_any_sync(0xffffffffu, x0 & 3));
        } else if (MODE == 9) {   // dependent FADD chain
            float f = __uint_as_float(x0);
            f = f + 1.0f; f = f + 1.0f; f = f + 1.0f; f = f + 1.0f;
            x0 = __float_as_uint(f);
        } else if (MODE == 10) {  // LDS dependent (pointer chase)
            x0 = my[x0 & 31];
        }
    }
    long long t1 = clock64();
    if (lane == 0) { out[2 * blockIdx.x] = (uint32_t)(t1 - t0); }
    out[1024 + threadIdx.x] = x0 + x1 + x2 + x3;
}

int main() {
    uint32_t* d;
    cudaMalloc(&d, 8192 * 4);
    uint32_t h[2];
    const char* names[] = {"butterfly3 (3 shfl+3 min+3 add)", "credux full", "4x credux per group + select", "smem STS/LDS min8",
                           "group-mask __reduce_min_sync", "vote.all", "shfl.idx", "ballot", "match_any", "4x fadd", "lds chase"};
#define RUN(M) \
    for (int ilp = 1; ilp <= 4; ilp *= 2) { \
        k<M><<<1, 32>>>(d, ilp); cudaDeviceSynchronize(); k<M><<<1, 32>>>(d, ilp); cudaDeviceSynchronize(); \
        cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost); \
        printf("%-36s ilp %d: %.1f cycles/iter\n", names[M], ilp, h[0] / (double)REP); }
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(10)
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
