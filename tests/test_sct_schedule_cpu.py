"""Host logic of the tile-per-block SC kernel (polarcodes_b200/csrc/decode_sct.cu): the pruned tree walk the host
compiles from a frozen pattern (op list), interpreted here in plain Python with the kernel's own data layout
(levels in half-split order, bit-packed partial-sum heap, 8-phase blocks), must reproduce the reference
decoder's decisions (SCDecoderMinsum.cpp:32-76 through the oracle).  No GPU needed."""
import ctypes as C

import numpy as np
import pytest

import polarcodes_b200 as pb

OP_TOP2, OP_LEVEL, OP_NODE8, OP_RATE0, OP_RATE1 = range(5)


def build_ops(n, mask):
    m = n.bit_length() - 1
    info = np.zeros(n // 32, np.uint32)
    fv = np.zeros(n // 32, np.uint32)
    for i, v in enumerate(mask):
        if v == -1:
            info[i >> 5] |= np.uint32(1 << (i & 31))
        elif v == 1:
            fv[i >> 5] |= np.uint32(1 << (i & 31))
    fn = pb.lib().pcb_dbg_sct_ops
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    cnt = fn(n, m, info.ctypes.data, fv.ctypes.data, None, 0)
    ops = np.zeros(cnt, np.uint32)
    assert fn(n, m, info.ctypes.data, fv.ctypes.data, ops.ctypes.data, cnt) == cnt
    return [int(o) for o in ops], info, fv


def f_ms(a, b):
    mn = min(abs(a), abs(b))
    return -mn if a * b < 0 else mn


def g_ms(a, b, c):
    return b - a if c else b + a


def sc_node(a, ib, fb):
    S = len(a)
    if S == 1:
        u = (0 if a[0] > 0 else 1) if (ib & 1) else (fb & 1)
        return u, u
    h = S // 2
    pL, uL = sc_node([f_ms(a[j], a[j + h]) for j in range(h)], ib, fb)
    pR, uR = sc_node([g_ms(a[j], a[j + h], (pL >> j) & 1) for j in range(h)], ib >> h, fb >> h)
    return (pL ^ pR) | (pR << h), uL | (uR << h)


def xform(x, bits):
    s = 1
    masks = {1: 0x55555555, 2: 0x33333333, 4: 0x0F0F0F0F, 8: 0x00FF00FF, 16: 0x0000FFFF}
    while s < bits:
        x ^= (x >> s) & masks[s]
        s *= 2
    return x


class Emu:
    def __init__(self, n, info, fv):
        self.n, self.m = n, n.bit_length() - 1
        self.info, self.fv = info, fv

    def cbit(self, pos):
        return (self.H[pos >> 5] >> (pos & 31)) & 1

    def finish_block(self, blk, ub, T):
        M = self.m
        self.uacc |= ub << (8 * (blk & 3))
        if (blk & 3) == 3:
            self.bits[blk >> 2] = self.uacc
            self.uacc = 0
        d2 = min(M, 3 + ((blk + 1) & -(blk + 1)).bit_length() - 1)
        if d2 == 3:
            self.w0 = (self.w0 & ~0xFF00 & 0xFFFFFFFF) | (T << 8)
            return
        T = (((self.w0 >> 8) ^ T) & 0xFF) | (T << 8)
        if d2 == 4:
            self.w0 = (self.w0 & 0xFFFF) | (T << 16)
            return
        T = (((self.w0 >> 16) ^ T) & 0xFFFF) | (T << 16)
        if d2 == M:
            return
        D = 1 << d2
        self.H[2 * D // 32 - 1] = T
        self.grow(32, D)

    def grow(self, w, D):
        while w < D:
            nw = w >> 5
            for kk in range(nw):
                self.H[(2 * D - 2 * w) // 32 + kk] = self.H[nw + kk] ^ self.H[(2 * D - w) // 32 + kk]
            w *= 2

    def finish_zero_words(self, blk0, nblk):
        for w in range(nblk // 4):
            self.bits[(blk0 >> 2) + w] = 0
        last = blk0 + nblk - 1
        d2 = min(self.m, 3 + ((last + 1) & -(last + 1)).bit_length() - 1)
        if d2 == self.m:
            return
        D, Dn = 1 << d2, nblk * 8
        for kk in range(Dn // 32):
            self.H[(2 * D - Dn) // 32 + kk] = 0
        self.grow(Dn, D)

    def decode(self, ops, y):
        n, m = self.n, self.m
        self.H = [0xDEADBEEF] * (n // 32)          # stale contents must never matter
        self.bits = [None] * (n // 32)
        self.w0, self.uacc = 0, 0
        L = {v: [None] * (n >> v) for v in range(2, m - 2)}
        bad = False
        for op in ops:
            typ, v, isg, fuse, aux = op & 15, (op >> 4) & 31, (op >> 9) & 1, (op >> 10) & 1, op >> 13
            if typ == OP_TOP2:
                g1, g2 = aux >> 1, aux & 1
                for r in range(n // 4):
                    x = int(format(r, f"0{m - 2}b")[::-1], 2)
                    y0, y1, y2, y3 = (float(t) for t in y[4 * r:4 * r + 4])
                    if g1:
                        la, lb = g_ms(y0, y1, self.cbit(n // 2 + x)), g_ms(y2, y3, self.cbit(n // 2 + x + n // 4))
                    else:
                        la, lb = f_ms(y0, y1), f_ms(y2, y3)
                    L[2][x] = g_ms(la, lb, self.cbit(n // 4 + x)) if g2 else f_ms(la, lb)
                if fuse:
                    for x in range(n // 8):
                        L[3][x] = f_ms(L[2][x], L[2][x + n // 8])
            elif typ == OP_LEVEL:
                S = n >> v
                for x in range(S):
                    c = 0
                    if isg:
                        c = self.cbit(S + x) if S >= 32 else (self.w0 >> (S + x)) & 1
                    a, b = L[v - 1][x], L[v - 1][x + S]
                    L[v][x] = g_ms(a, b, c) if isg else f_ms(a, b)
                if fuse:
                    for x in range(S // 2):
                        L[v + 1][x] = f_ms(L[v][x], L[v][x + S // 2])
            elif typ == OP_NODE8:
                ib = (int(self.info[aux >> 2]) >> (8 * (aux & 3))) & 0xFF
                fb = (int(self.fv[aux >> 2]) >> (8 * (aux & 3))) & 0xFF
                T, ub = sc_node(list(L[m - 3]), ib, fb)
                self.finish_block(aux, ub, T)
            elif typ == OP_RATE0:
                nblk = 1 << v
                if nblk >= 4:
                    self.finish_zero_words(aux, nblk)
                else:
                    for b in range(nblk):
                        self.finish_block(aux + b, 0, 0)
            elif typ == OP_RATE1:
                Sz = 1 << v
                src = L[m - v]
                beta = 0
                for j in range(Sz):
                    bad = bad or not (abs(src[j]) >= 2.0 ** -511)
                    beta |= (0 if src[j] > 0 else 1) << j
                u = xform(beta, Sz)
                for b in range(Sz // 8):
                    ub = (u >> (8 * b)) & 0xFF
                    self.finish_block(aux + b, ub, xform(ub, 8))
            else:
                raise AssertionError(f"unknown op {op:#x}")
        u = np.zeros(n, np.int32)
        for i in range(n):
            u[i] = (self.bits[i >> 5] >> (i & 31)) & 1
        return u, bad


@pytest.mark.parametrize("n,k,snr", [(128, 64, 2.0), (256, 128, 1.0), (1024, 512, 2.0), (1024, 820, 4.0), (512, 100, 0.0)])
def test_op_list_reproduces_the_reference_decoder(oracle, n, k, snr):
    code = pb.Code(n, k, 0, snr)
    ops, info, fv = build_ops(n, code.mask)
    kinds = {o & 15 for o in ops}
    assert OP_TOP2 in kinds and OP_NODE8 in kinds
    rng = np.random.default_rng(n + k)
    B = 6
    y = 1.0 + code.sigma * rng.standard_normal((B, n))       # all-zero codeword, BPSK + AWGN
    y[1] = np.round(y[1] * 4) / 4                             # quantised values: exact ties and zeros inside the tree
    want = oracle.decode(2, n, 1, k, 0, code.sigma, y, code.mask)
    emu = Emu(n, info, fv)
    nbad = 0
    for b in range(B):
        got, bad = emu.decode(ops, y[b])
        nbad += bad
        if not bad:                                           # a flagged frame is decoded again by the generic kernel
            assert (got == want[b]).all(), f"frame {b}: {np.flatnonzero(got != want[b])[:8]}"
    assert nbad < B


def test_op_list_with_frozen_ones_keeps_those_subtrees(oracle):
    n, k = 256, 128
    code = pb.Code(n, k, 0, 2.0)
    mask = np.array(code.mask).copy()
    frozen = np.flatnonzero(mask == 0)
    mask[frozen[::3]] = 1                                     # frozen values 1: not prunable as all-zero subtrees
    ops, info, fv = build_ops(n, mask)
    rng = np.random.default_rng(5)
    y = 1.0 + code.sigma * rng.standard_normal((4, n))
    want = oracle.decode(2, n, 1, k, 0, code.sigma, y, mask)
    emu = Emu(n, info, fv)
    for b in range(4):
        got, bad = emu.decode(ops, y[b])
        assert bad or (got == want[b]).all()
