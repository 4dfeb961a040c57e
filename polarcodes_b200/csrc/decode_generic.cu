// Generic batched SC / SCL decoder: ONE WARP DECODES ONE FRAME, any n = 2^m, any list size
// L in {1,2,4,8,16,32}.  This is the reference-exact path for every configuration; the
// list-8 hot path has a tuned sibling (decode_scl8.cu) that must agree with it bit for bit.
//
// What is reproduced from the reference (paths relative to /root/reference/PolarCodes/):
//   * LLR schedule  iterativelyCalcP, SCLDecoderMinsum.cpp:225-276: at phase i recompute layers
//     m-d..m, d = min(m-1, ctz(i)) (i = 0: m-1); the first layer uses g iff bit d of i is set,
//     deeper layers use f; half-split pairing (j, j+s).
//   * partial sums  iterativelyUpdateC, :278-306 / getArrayPointerCW :202-223.
//   * fork / prune  continuePaths_UnfrozenBit, :315-400, including the slot discipline of
//     the LIFO free stack (:133-168) and the median tie rule (:346-371).
//   * final pick    :432-473 (slots in index order, first CRC pass wins, else slot L-1).
//   * SC decoders   SCDecoderMinsum.cpp:32-76 / SCDecoder.cpp:38-85 run through the same
//     schedule with L = 1 on bit-reversed inputs: adjacent pairing on y equals half-split
//     pairing on y[bitrev(j)], every LLR is produced by the same operations on the same
//     operands, and decisions come out in the same order.
//
// B200 mapping:
//   * lazy copy without a pool: at a given phase EVERY live path rewrites the same layers,
//     so slot p always writes its own arrays and only needs, per layer, the slot it must
//     READ from (a log2(L)-bit field per layer packed in one 64-bit word per path).  A clone
//     copies that word (one shuffle) instead of arrays; nothing is ever freed.
//   * partial sums are bit-packed in a per-slot "heap" of 2n bits: the layer of size S sits at
//     bits [S, 2S).  Layers below 32 bits share word 0, which lives in a register of lane p.
//   * prune: each candidate's strict rank among the 2L candidates is counted with all-pairs
//     integer compares on the metric bit patterns (metrics are non-negative floats), ties at
//     the median fall to a rare exact path; survivors are then renumbered with ballots.
#include "decode_common.cuh"
#include "decode_launch.h"

namespace pcb {

template <typename T, int L, int MODE>
__global__ void __launch_bounds__(256) decode_generic_kernel(const DecodeArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ unsigned short crc_tab[256];
    constexpr int LB = Log2<L>::value;
    constexpr int FB = LB > 0 ? LB : 1;
    constexpr unsigned FULL = 0xffffffffu;
    typedef unsigned long long u64;

    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const int n = A.n, m = A.m;
    const int ls = A.first_smem_layer;

    // CRC table (Resources.cpp:29-62 poly 0x31 / :73-106 poly 0x1021), built once per block
    if (MODE == MODE_SCL && A.crc) {
        for (int t = threadIdx.x; t < 256; t += blockDim.x) {
            unsigned r = A.crc == 8 ? (unsigned)t : ((unsigned)t << 8);
            for (int b = 0; b < 8; b++) {
                if (A.crc == 8) r = (r & 0x80) ? ((r << 1) ^ 0x31) & 0xFF : (r << 1) & 0xFF;
                else r = (r & 0x8000) ? ((r << 1) ^ 0x1021) & 0xFFFF : (r << 1) & 0xFFFF;
            }
            crc_tab[t] = (unsigned short)r;
        }
    }
    __syncthreads();

    // ---- per-warp shared-memory slice -------------------------------------------------
    const int HW = (2 * n >= 32) ? (2 * n / 32) : 1;           // heap words per slot
    const int EW = (A.stride + 31) >> 5;                        // packed solution words
    unsigned char* ws = smem_raw + (size_t)wib * A.warp_smem_bytes;
    const WarpLayout lay = warp_layout((int)sizeof(T), L, n, ls, A.stride, MODE);
    T* Ps = reinterpret_cast<T*>(ws + lay.off_P);
    uint32_t* H = reinterpret_cast<uint32_t*>(ws + lay.off_H);
    u64* pidx_s = reinterpret_cast<u64*>(ws + lay.off_pidx);
    u64* cidx_s = reinterpret_cast<u64*>(ws + lay.off_cidx);
    uint32_t* keys = reinterpret_cast<uint32_t*>(ws + lay.off_keys);
    uint32_t* E = reinterpret_cast<uint32_t*>(ws + lay.off_E);
    const int gwarp = blockIdx.x * wpb + wib;
    const int total_warps = gridDim.x * wpb;
    T* Pg = reinterpret_cast<T*>(A.scratch + (size_t)gwarp * A.scratch_per_warp);

    auto Pl = [&](int l) -> T* {  // layer l >= 1, slot stride n>>l
        return (l >= ls) ? Ps + (size_t)L * ((n >> (ls - 1)) - (n >> (l - 1)))
                         : Pg + (size_t)L * (n - (n >> (l - 1)));
    };
    auto field = [&](u64 w, int layer) -> int { return L == 1 ? 0 : (int)((w >> (FB * (layer - 1))) & (L - 1)); };

    u64 REP = 0;
    if (L > 1)
        for (int t = 0; t < 64 / FB; t++) REP |= 1ull << (FB * t);

    const double inv_s2 = A.sigma * A.sigma;

    const int nframes = A.use_redo_list ? *A.redo_count : A.B;
    for (int fidx = gwarp; fidx < nframes; fidx += total_warps) {
        const int frame = A.use_redo_list ? A.redo_list[fidx] : fidx;
        // a batch may be floats for the frames below y64_first and doubles behind (host pipeline of the list decoder)
        const bool is_f32 = A.y_f32 && !(A.y64_tail && frame >= A.y64_first);
        const float* yf = reinterpret_cast<const float*>(A.y) + (size_t)frame * n;
        const double* yd = A.y64_tail ? A.y64_tail + (size_t)(frame - A.y64_first) * n
                                      : reinterpret_cast<const double*>(A.y) + (size_t)frame * n;
        auto ch = [&](int j) -> T {
            const int idx = (MODE == MODE_SCL) ? j : (int)(__brev((unsigned)j) >> (32 - m));
            T v = is_f32 ? (T)yf[idx] : (T)yd[idx];
            if (MODE == MODE_SC_EXACT) {  // SCDecoder.cpp:15-17, :77-79
                v = (T)((double)v * 2. / inv_s2);
                if (isinf((double)v)) v = (T)DBL_MAX;
            }
            return v;
        };

        int nact = 1;              // live slots are [L - nact, L)  (growth pops L-1, L-2, ...)
        float met = 0.f;           // lane p: path metric of slot p
        u64 pidx = 0, cidx = 0;    // lane p: per-layer source slot of the LLR / partial-sum arrays
        uint32_t w0 = 0;           // lane p: heap word 0 (partial-sum layers of < 32 bits)
        uint32_t uacc = 0, iw = 0, fw = 0;
        if (lane < L) { H[lane * HW] = 0; pidx_s[lane] = 0; cidx_s[lane] = 0; }
        __syncwarp();

        for (int i = 0; i < n; ++i) {
            const int d = i ? min(m - 1, __ffs(i) - 1) : m - 1;
            const int ll = m - d;
            const bool gfirst = (i >> d) & 1;
            const int pbase = L - nact;
            if ((i & 31) == 0) { iw = __ldg(A.info_mask + (i >> 5)); fw = __ldg(A.frozen_val + (i >> 5)); }

            // ---- LLR layers ll..m ------------------------------------------------------
            for (int l = ll; l <= m; ++l) {
                const int lg = m - l, s = 1 << lg;
                const bool first = (l == ll);
                const int total = nact << lg;
                T* dst = Pl(l);
                const T* srcb = (l > 1) ? Pl(l - 1) : nullptr;
                for (int e = lane; e < total; e += 32) {
                    const int p = pbase + (e >> lg), j = e & (s - 1);
                    T a, b;
                    if (l == 1) { a = ch(j); b = ch(j + s); }
                    else {
                        const int ss = first ? field(pidx_s[p], l - 1) : p;
                        const T* sp = srcb + (size_t)ss * 2 * s;
                        a = sp[j]; b = sp[j + s];
                    }
                    T o;
                    if (first && gfirst) {
                        int c;
                        if (s < 32) c = (H[p * HW] >> (s + j)) & 1;
                        else c = (H[field(cidx_s[p], l) * HW + (s >> 5) + (j >> 5)] >> (j & 31)) & 1;
                        o = (MODE == MODE_SC_EXACT) ? (T)g_exact((double)a, (double)b, c) : g_minsum<T>(a, b, c);
                    } else {
                        o = (MODE == MODE_SC_EXACT) ? (T)f_exact((double)a, (double)b) : f_minsum<T>(a, b);
                    }
                    dst[(size_t)p * s + j] = o;
                }
                __syncwarp();
            }
            const bool active = lane >= pbase && lane < L;
            if (active && L > 1) {  // layers >= ll are now this slot's own
                const int sh = FB * (ll - 1);
                const u64 low = sh >= 64 ? ~0ull : ((1ull << sh) - 1);
                pidx = (pidx & low) | ((REP * (u64)lane) & ~low);
            }
            T llr = (T)0;
            if (active) llr = Pl(m)[lane];
            const bool isinfo = (iw >> (i & 31)) & 1;
            const int fv = (fw >> (i & 31)) & 1;
            int u = fv;

            if (MODE != MODE_SCL) {
                // SCDecoderMinsum.cpp:33-47: decision `> 0 ? 0 : 1` (an LLR of exactly 0 decodes to 1)
                if (isinfo) u = (llr > (T)0) ? 0 : 1;
                uacc |= (uint32_t)u << (i & 31);
                if ((i & 31) == 31 || i == n - 1) { if (lane == 0) E[i >> 5] = uacc; uacc = 0; }
            } else if (!isinfo) {
                if (active) met += pm_penalty((float)llr, fv);   // :413-418
            } else {
                const float pf0 = met + pm_penalty((float)llr, 0);   // :319-328
                const float pf1 = met + pm_penalty((float)llr, 1);
                if (nact < L) {
                    // growth (:329-341): parent p clones into q = 2*pbase-1-p (LIFO pops)
                    const bool isclone = lane >= pbase - nact && lane < pbase;
                    const int src = isclone ? (2 * pbase - 1 - lane) : lane;
                    const float m1 = __shfl_sync(FULL, pf1, src);
                    const u64 pi = __shfl_sync(FULL, pidx, src);
                    const u64 ci = __shfl_sync(FULL, cidx, src);
                    const uint32_t ww = __shfl_sync(FULL, w0, src);
                    if (isclone) { met = m1; pidx = pi; cidx = ci; w0 = ww; u = 1; }
                    else if (active) { met = pf0; u = 0; }
                    nact *= 2;
                } else {
                    // prune (:342-399)
                    constexpr int C = 2 * L;
                    if (lane < L) { keys[2 * lane] = __float_as_uint(pf0); keys[2 * lane + 1] = __float_as_uint(pf1); }
                    __syncwarp();
                    u64 keepmask;
                    uint32_t key0, key1 = 0;
                    if constexpr (C <= 32) {
                        constexpr int PARTS = (32 / C < C) ? 32 / C : C;
                        constexpr int JN = C / PARTS;
                        const int c = lane % C, part = (lane / C) % PARTS;
                        key0 = keys[c];
                        int cnt = 0;
#pragma unroll
                        for (int jj = 0; jj < JN; jj++) cnt += keys[part * JN + jj] < key0;
#pragma unroll
                        for (int off = C; off < C * PARTS; off <<= 1) cnt += __shfl_xor_sync(FULL, cnt, off);
                        const unsigned maskC = C == 32 ? FULL : ((1u << (C & 31)) - 1);
                        unsigned km = __ballot_sync(FULL, cnt < L) & maskC;
                        if (__popc(km) != L) {  // a tie straddles the median: pass 2, :362-371
                            const unsigned in0 = (km >> c) & 1;
                            const unsigned tkey = __reduce_max_sync(FULL, in0 ? key0 : 0u);
                            const unsigned tied = __ballot_sync(FULL, in0 && key0 == tkey) & maskC;
                            const int need = L - (__popc(km) - __popc(tied));
                            const bool kp = in0 && (key0 < tkey || __popc(tied & ((1u << c) - 1)) < need);
                            km = __ballot_sync(FULL, kp) & maskC;
                        }
                        keepmask = km;
                    } else {
                        key0 = keys[lane]; key1 = keys[lane + 32];
                        int c0 = 0, c1 = 0;
                        for (int j = 0; j < C; j++) { const uint32_t kj = keys[j]; c0 += kj < key0; c1 += kj < key1; }
                        unsigned k0m = __ballot_sync(FULL, c0 < L), k1m = __ballot_sync(FULL, c1 < L);
                        if (__popc(k0m) + __popc(k1m) != L) {
                            const unsigned in0 = (k0m >> lane) & 1, in1 = (k1m >> lane) & 1;
                            unsigned v = in0 ? key0 : 0u;
                            if (in1 && key1 > v) v = key1;
                            const unsigned tkey = __reduce_max_sync(FULL, v);
                            const unsigned t0m = __ballot_sync(FULL, in0 && key0 == tkey);
                            const unsigned t1m = __ballot_sync(FULL, in1 && key1 == tkey);
                            const int need = L - (__popc(k0m) + __popc(k1m) - __popc(t0m) - __popc(t1m));
                            const unsigned lt = lanemask_lt();
                            const bool kp0 = in0 && (key0 < tkey || __popc(t0m & lt) < need);
                            const bool kp1 = in1 && (key1 < tkey || __popc(t0m) + __popc(t1m & lt) < need);
                            k0m = __ballot_sync(FULL, kp0); k1m = __ballot_sync(FULL, kp1);
                        }
                        keepmask = (u64)k0m | ((u64)k1m << 32);
                    }
                    const int k0 = lane < L ? (int)((keepmask >> (2 * lane)) & 1) : 0;
                    const int k1 = lane < L ? (int)((keepmask >> (2 * lane + 1)) & 1) : 0;
                    const unsigned nonem = __ballot_sync(FULL, lane < L && !k0 && !k1);
                    const unsigned bothm = __ballot_sync(FULL, lane < L && k0 && k1);
                    const bool none = (nonem >> lane) & 1;
                    int src = lane;
                    if (nonem) {
                        // killed slots are pushed ascending (:372-376), cloning parents pop in
                        // ascending order (:382-389): the j-th killed slot (ascending) receives
                        // the clone of the j-th HIGHEST cloning parent.
                        if (none) {
                            const int jr = __popc(nonem & lanemask_lt());
                            unsigned bm = bothm;
                            for (int t = 0; t < jr; t++) bm &= ~(0x80000000u >> __clz(bm));
                            src = 31 - __clz(bm);
                        }
                        const float m1 = __shfl_sync(FULL, pf1, src);
                        const u64 pi = __shfl_sync(FULL, pidx, src);
                        const u64 ci = __shfl_sync(FULL, cidx, src);
                        const uint32_t ww = __shfl_sync(FULL, w0, src);
                        if (none) { met = m1; pidx = pi; cidx = ci; w0 = ww; u = 1; }
                    }
                    if (!none && lane < L) {
                        if (k0) { met = pf0; u = 0; } else { met = pf1; u = 1; }
                    }
                }
            }

            // ---- partial sums: write bit u of every live slot, combine finished subtrees ----
            {
                const int pb = L - nact;
                const bool act = lane >= pb && lane < L;
                if (!(i & 1)) {
                    if (act) w0 = (w0 & ~2u) | ((uint32_t)u << 1);
                } else {
                    const int d2 = min(m, __ffs(i + 1) - 1);
                    const int D = 1 << d2;
                    uint32_t Tt = (uint32_t)u;
                    if (act)
                        for (int S = 1; S < D && S < 32; S <<= 1) Tt = (((w0 >> S) ^ Tt) & ((1u << S) - 1)) | (Tt << S);
                    if (D < 32) {
                        if (act) w0 = (w0 & ~(((1u << D) - 1) << D)) | (Tt << D);
                    } else {
                        if (act) { H[lane * HW + 2 * D / 32 - 1] = Tt; cidx_s[lane] = cidx; }
                        __syncwarp();
                        for (int w = 32; w < D; w <<= 1) {
                            const int lgw = 31 - __clz(w);
                            const int layc = m - lgw, nw = w >> 5, lgn = lgw - 5;
                            const int total = nact << lgn;
                            for (int e = lane; e < total; e += 32) {
                                const int p = pb + (e >> lgn), kk = e & (nw - 1);
                                const int cs = field(cidx_s[p], layc);
                                H[p * HW + (2 * D - 2 * w) / 32 + kk] =
                                    H[cs * HW + nw + kk] ^ H[p * HW + (2 * D - w) / 32 + kk];
                            }
                            __syncwarp();
                        }
                        if (act && d2 < m && L > 1) {
                            const int sh = FB * (m - d2 - 1);
                            cidx = (cidx & ~((u64)(L - 1) << sh)) | ((u64)lane << sh);
                        }
                    }
                }
                if (act) { H[lane * HW] = w0; pidx_s[lane] = pidx; cidx_s[lane] = cidx; }
                __syncwarp();
            }
        }  // phases

        // ---- output -------------------------------------------------------------------
        if (MODE != MODE_SCL) {
            __syncwarp();
            if (A.sol)
                for (int t = lane; t < n; t += 32) A.sol[(size_t)frame * n + t] = (E[t >> 5] >> (t & 31)) & 1;
            if (A.sol_bits)
                for (int t = lane; t < EW; t += 32) A.sol_bits[(size_t)frame * EW + t] = E[t];
            __syncwarp();
            continue;
        }
        // SCL: u-hat = x-hat F^(x)m per slot (natural order, :434-435), in place on heap layer 0
        int xoff, NW;
        if (n < 32) {
            if (lane < L) {
                uint32_t xb = (w0 >> n) & ((1u << n) - 1);
                if (n > 1) xb ^= (xb >> 1) & 0x55555555u;
                if (n > 2) xb ^= (xb >> 2) & 0x33333333u;
                if (n > 4) xb ^= (xb >> 4) & 0x0F0F0F0Fu;
                if (n > 8) xb ^= (xb >> 8) & 0x00FF00FFu;
                H[lane * HW] = xb;
            }
            xoff = 0; NW = 1;
            __syncwarp();
        } else {
            xoff = n / 32; NW = n / 32;
            const int lgN = m - 5;
            for (int e = lane; e < (L << lgN); e += 32) {
                const int p = e >> lgN, a = e & (NW - 1);
                uint32_t x = H[p * HW + xoff + a];
                x ^= (x >> 1) & 0x55555555u;
                x ^= (x >> 2) & 0x33333333u;
                x ^= (x >> 4) & 0x0F0F0F0Fu;
                x ^= (x >> 8) & 0x00FF00FFu;
                x ^= (x >> 16) & 0x0000FFFFu;
                H[p * HW + xoff + a] = x;
            }
            __syncwarp();
            for (int sw = 1; sw < NW; sw <<= 1) {
                for (int e = lane; e < (L << lgN); e += 32) {
                    const int p = e >> lgN, a = e & (NW - 1);
                    if (!(a & sw)) H[p * HW + xoff + a] ^= H[p * HW + xoff + a + sw];
                }
                __syncwarp();
            }
        }
        // info (+crc) bits of every slot, packed LSB-first (:436-441)
        for (int p = 0; p < L; p++) {
            for (int t0 = 0; t0 < A.stride; t0 += 32) {
                const int t = t0 + lane;
                int bit = 0;
                if (t < A.stride) { const int pos = A.info_pos[t]; bit = (H[p * HW + xoff + (pos >> 5)] >> (pos & 31)) & 1; }
                const unsigned wd = __ballot_sync(FULL, bit);
                if (lane == 0) E[p * EW + (t0 >> 5)] = wd;
            }
        }
        __syncwarp();
        // CRC of every slot in parallel (:442-472); first pass in slot order wins
        bool ok = false;
        if (A.crc && lane < L) {
            const uint32_t* e = E + lane * EW;
            unsigned r = A.crc == 8 ? 0xFFu : 0xFFFFu;
            for (int b = 0; b < A.k / 8; b++) {
                const unsigned byte = __brev((e[b >> 2] >> (8 * (b & 3))) & 0xFFu) >> 24;
                if (A.crc == 8) r = crc_tab[(r ^ byte) & 0xFF];
                else r = ((r << 8) ^ crc_tab[((r >> 8) ^ byte) & 0xFF]) & 0xFFFF;
            }
            unsigned rx = 0;
            for (int t = 0; t < A.crc; t++) {
                const int q = A.k + t;
                rx = (rx << 1) | ((e[q >> 5] >> (q & 31)) & 1);
            }
            ok = (rx == r);
        }
        const unsigned okm = __ballot_sync(FULL, ok);
        const int chosen = okm ? (__ffs(okm) - 1) : (L - 1);
        if (A.sol)
            for (int t = lane; t < A.stride; t += 32)
                A.sol[(size_t)frame * A.stride + t] = (E[chosen * EW + (t >> 5)] >> (t & 31)) & 1;
        if (A.sol_bits)
            for (int t = lane; t < EW; t += 32) A.sol_bits[(size_t)frame * EW + t] = E[chosen * EW + t];
        __syncwarp();
    }
}

template <typename T, int L, int MODE>
static cudaError_t launch_one(const DecodeArgs& a, int blocks, int wpb, cudaStream_t st) {
    const size_t smem = (size_t)a.warp_smem_bytes * wpb;
    cudaError_t e = cudaFuncSetAttribute(decode_generic_kernel<T, L, MODE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    decode_generic_kernel<T, L, MODE><<<blocks, wpb * 32, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_decode_generic(const DecodeArgs& a, int L, int mode, int fp64, int blocks, int wpb,
                                  cudaStream_t st) {
    if (mode == MODE_SC_EXACT) return launch_one<double, 1, MODE_SC_EXACT>(a, blocks, wpb, st);
    if (mode == MODE_SC_MINSUM)
        return fp64 ? launch_one<double, 1, MODE_SC_MINSUM>(a, blocks, wpb, st)
                    : launch_one<float, 1, MODE_SC_MINSUM>(a, blocks, wpb, st);
    switch (L) {
        case 1: return launch_one<float, 1, MODE_SCL>(a, blocks, wpb, st);
        case 2: return launch_one<float, 2, MODE_SCL>(a, blocks, wpb, st);
        case 4: return launch_one<float, 4, MODE_SCL>(a, blocks, wpb, st);
        case 8: return launch_one<float, 8, MODE_SCL>(a, blocks, wpb, st);
        case 16: return launch_one<float, 16, MODE_SCL>(a, blocks, wpb, st);
        case 32: return launch_one<float, 32, MODE_SCL>(a, blocks, wpb, st);
    }
    return cudaErrorInvalidValue;
}


// ---- no-allocation variant (pcb_decode_batch_dev_noalloc) --------------------------------------
// The large LLR layers of this kernel live in a global scratch (DecodeArgs::scratch, one slice per
// resident warp).  For callers that may not allocate (CUDA graph capture, serving loops), the same
// kernel runs with its slices in this file-scope array: every GPU that loads the library holds
// PCB_NOALLOC_SCRATCH_BYTES of it until the process exits.  Nothing in it has to start at zero (every
// layer is written before it is read within a frame); all calls on a device share it, so two calls
// that overlap on the GPU race.
static __device__ __align__(256) unsigned char g_noalloc_scratch[PCB_NOALLOC_SCRATCH_BYTES];

// The array's name is a host-side stand-in in host code: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current (called from pcb_create).
cudaError_t noalloc_scratch_address(unsigned char** ptr, size_t* bytes) {
    static unsigned char* cached[64] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!cached[dev]) {
        unsigned char* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        cached[dev] = p;
    }
    *ptr = cached[dev];
    if (bytes) *bytes = PCB_NOALLOC_SCRATCH_BYTES;
    return cudaSuccess;
}


template <typename T, int L, int MODE>
static cudaError_t occ_one(int wpb, size_t smem, int* bps) {
    cudaError_t e = cudaFuncSetAttribute(decode_generic_kernel<T, L, MODE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, decode_generic_kernel<T, L, MODE>, wpb * 32, smem);
}

cudaError_t decode_generic_occupancy(int L, int mode, int fp64, int wpb, size_t smem, int* bps) {
    if (mode == MODE_SC_EXACT) return occ_one<double, 1, MODE_SC_EXACT>(wpb, smem, bps);
    if (mode == MODE_SC_MINSUM)
        return fp64 ? occ_one<double, 1, MODE_SC_MINSUM>(wpb, smem, bps) : occ_one<float, 1, MODE_SC_MINSUM>(wpb, smem, bps);
    switch (L) {
        case 1: return occ_one<float, 1, MODE_SCL>(wpb, smem, bps);
        case 2: return occ_one<float, 2, MODE_SCL>(wpb, smem, bps);
        case 4: return occ_one<float, 4, MODE_SCL>(wpb, smem, bps);
        case 8: return occ_one<float, 8, MODE_SCL>(wpb, smem, bps);
        case 16: return occ_one<float, 16, MODE_SCL>(wpb, smem, bps);
        case 32: return occ_one<float, 32, MODE_SCL>(wpb, smem, bps);
    }
    return cudaErrorInvalidValue;
}

}  // namespace pcb
