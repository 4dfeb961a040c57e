// Shared by the tile-per-block SC kernels (decode_sct.cu, decode_sc8.cu): the op list the host compiles from a frozen
// pattern (pruned tree walk), and the leaf-walk state of a warp (decisions, bit-packed partial sums and their combine).
#pragma once
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#ifndef PCB_T_U
#define PCB_T_U 4          // rows per loop iteration of a level op (loads in flight: 2U, fused 4U)
#endif
#ifndef PCB_T_PARMIN
#define PCB_T_PARMIN 4     // a level op is split between the warps when each gets at least this many rows (a multiple of PCB_T_U)
#endif

namespace pcb {
namespace sct {

enum : uint32_t { OP_TOP2 = 0, OP_LEVEL = 1, OP_NODE8 = 2, OP_RATE0 = 3, OP_RATE1 = 4 };
// op word: type [0,4) | v [4,9) | isg 9 | fuse 10 | parallel 11 | sync_before 12 | aux [13,32)
static inline uint32_t mk(uint32_t type, int v, int isg, int fuse, int par, int aux) {
    return type | ((uint32_t)v << 4) | ((uint32_t)isg << 9) | ((uint32_t)fuse << 10) | ((uint32_t)par << 11) | ((uint32_t)aux << 13);
}

__device__ __forceinline__ uint32_t xform8(uint32_t x) {       // u <-> partial sums of a size-8 node, packed
    x ^= (x >> 1) & 0x55u; x ^= (x >> 2) & 0x33u; x ^= (x >> 4) & 0x0Fu;
    return x;
}

// State of warp 0 (the one that walks the leaves): one frame per lane.
struct Walk {
    uint32_t w0, uacc;       // partial sums of sizes < 32 (size S at bits [S, 2S)); decisions of the current 32 phases
    uint32_t* H;             // this lane's column of the heap: word w at H[w * 32]
    uint32_t* bits;          // this lane's packed decisions
    bool store;              // frame < B
    int M;
    // one 8-phase block is decided: ub = its decisions in phase order, Tps = its partial sums
    __device__ __forceinline__ void finish_block(int blk, uint32_t ub, uint32_t Tps) {
        uacc |= ub << (8 * (blk & 3));
        if ((blk & 3) == 3) {
            if (store) bits[blk >> 2] = uacc;
            uacc = 0;
        }
        const int d2 = min(M, 3 + __ffs(blk + 1) - 1);     // SCLDecoderMinsum.cpp:287-305
        if (d2 == 3) { w0 = (w0 & ~0xFF00u) | (Tps << 8); return; }
        Tps = (((w0 >> 8) ^ Tps) & 0xFFu) | (Tps << 8);
        if (d2 == 4) { w0 = (w0 & 0x0000FFFFu) | (Tps << 16); return; }
        Tps = (((w0 >> 16) ^ Tps) & 0xFFFFu) | (Tps << 16);
        if (d2 == M) return;                               // the whole codeword: nobody reads it
        const int D = 1 << d2;
        H[(2 * D / 32 - 1) * 32] = Tps;
        grow(32, D);
    }
    // the last `w` bits of the size-D array at heap bits [D, 2D) are in place: prepend (left node ^ them) until it is full
    __device__ __forceinline__ void grow(int w, int D) {
#pragma unroll 1
        for (; w < D; w <<= 1) {
            const int nw = w >> 5;
#pragma unroll 1
            for (int kk = 0; kk < nw; kk++)
                H[((2 * D - 2 * w) / 32 + kk) * 32] = H[(nw + kk) * 32] ^ H[((2 * D - w) / 32 + kk) * 32];
        }
    }
    // an all-frozen node of nblk >= 4 blocks (a multiple of 4, aligned): zero decisions, zero partial sums
    __device__ __forceinline__ void finish_zero_words(int blk0, int nblk) {
        if (store) {
#pragma unroll 1
            for (int w = 0; w < nblk / 4; w++) bits[(blk0 >> 2) + w] = 0u;
        }
        const int last = blk0 + nblk - 1;
        const int d2 = min(M, 3 + __ffs(last + 1) - 1);
        if (d2 == M) return;
        const int D = 1 << d2, Dn = nblk * 8;
#pragma unroll 1
        for (int kk = 0; kk < Dn / 32; kk++) H[((2 * D - Dn) / 32 + kk) * 32] = 0u;
        grow(Dn, D);
    }
};

// ---- host: compile the pruned tree walk of a frozen pattern into the op list --------------------------------
struct Builder {
    int n, m, W;
    const uint32_t* info;    // n/32 words, bit set = information position
    const uint32_t* fv;      // frozen values
    std::vector<uint32_t> ops, pops;
    int count(const uint32_t* w, int lo, int S) const {
        int c = 0;
        for (int i = lo; i < lo + S; i++) c += (w[i >> 5] >> (i & 31)) & 1u;
        return c;
    }
    bool rate0(int lo, int S) const { return S >= 8 && count(info, lo, S) == 0 && count(fv, lo, S) == 0; }
    bool use_rate1 = true;   // all-information subtrees decided from the signs of their inputs (needs the redo path)
    bool rate1(int lo, int S) const { return use_rate1 && S >= 8 && S <= 32 && count(info, lo, S) == S; }
    void node(int lo, int d) {       // a node at depth d: its input is level d (pruned from depth 1 on)
        const int S = n >> d;
        if (d >= 1 && rate0(lo, S)) { int lg = 0; while ((8 << lg) < S) lg++; ops.push_back(mk(OP_RATE0, lg, 0, 0, 0, lo / 8)); return; }
        if (d >= 1 && rate1(lo, S)) { int lg = 3; while ((1 << lg) < S) lg++; ops.push_back(mk(OP_RATE1, lg, 0, 0, 0, lo / 8)); return; }
        if (S == 8) { ops.push_back(mk(OP_NODE8, 0, 0, 0, 0, lo / 8)); return; }
        const int h = S / 2;
        if (!rate0(lo, h)) ops.push_back(mk(OP_LEVEL, d + 1, 0, 0, 0, 0));
        node(lo, d + 1);
        if (!rate0(lo + h, h)) ops.push_back(mk(OP_LEVEL, d + 1, 1, 0, 0, 0));
        node(lo + h, d + 1);
    }
    // the walk from the root with one op per level, nothing fused, everything executed by one warp (decode_sc8.cu)
    void build_flat() {
        node(0, 0);
    }
    void build() {
        for (int q = 0; q < 4; q++) {
            const int lo = q * (n / 4);
            if (!rate0(lo, n / 4)) ops.push_back(mk(OP_TOP2, 2, 0, 0, 1, q));
            node(lo, 2);
        }
        // fuse an op with the f-level that follows it; mark what all warps execute
        std::vector<uint32_t> out;
        for (size_t i = 0; i < ops.size(); i++) {
            uint32_t op = ops[i];
            const uint32_t type = op & 15u;
            const int v = (op >> 4) & 31;
            if ((type == OP_TOP2 || type == OP_LEVEL) && i + 1 < ops.size()) {
                const uint32_t nx = ops[i + 1];
                if ((nx & 15u) == OP_LEVEL && (int)((nx >> 4) & 31) == v + 1 && !((nx >> 9) & 1u) && v + 1 <= m - 3) { op |= 1u << 10; i++; }
            }
            if (type == OP_LEVEL) {
                // all warps: the heap must hold the partial sums (S >= 32; warp 0 keeps the smaller ones in a register)
                // and every warp gets at least one full iteration
                const int S = n >> v;
                const bool fuse = (op >> 10) & 1u;
                if (S >= 32 && (fuse ? S / 2 : S) / W >= PCB_T_PARMIN) op |= 1u << 11;
            }
            out.push_back(op);
        }
        // a barrier in front of an op all warps execute when warp 0 has worked alone since the last one
        bool serial_since = true;
        for (auto& op : out) {
            const uint32_t type = op & 15u;
            const bool par = type == OP_TOP2 || (type == OP_LEVEL && ((op >> 11) & 1u));
            if (par) { if (serial_since) op |= 1u << 12; serial_since = false; pops.push_back(op); }
            else serial_since = true;
        }
        ops.swap(out);
    }
};

}  // namespace sct
}  // namespace pcb
