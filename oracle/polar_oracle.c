/* TEST INFRASTRUCTURE ONLY — see polar_oracle.h.
 *
 * Plain-C restatement of the hot path of dim49v/polarCodes.  File:line citations refer to
 * /root/reference/PolarCodes/.  The list decoder is restated as an EAGER-copy model (every
 * path slot owns its own LLR / partial-sum arrays); the reference's lazy-copy array pool
 * (SCLDecoderMinsum.cpp:133-223) never copies contents and is unobservable — what IS
 * observable, and is reproduced exactly, is the path-slot discipline: LIFO free stack,
 * ascending kills, ascending clones, median ties broken in slot order, final pick in slot
 * order.  Build with -ffp-contract=off (no FMA contraction anywhere).
 */
#include "polar_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ construction ---- */

/* PolarCodes.cpp:123 — Eb/N0 with rate k/n (crc bits not counted). */
double po_sigma(int n, int k, double snr_db) { return sqrt(n / (2. * k * pow(10, snr_db / 10))); }

/* PolarCodes.cpp:31-44 */
static double awgn_phi(double x) {
    if (x > 12) return 0.9861 * x - 2.3152;
    if (x > 3.5) return x * (0.009005 * x + 0.7694) - 0.9507;
    if (x > 1) return x * (0.062883 * x + 0.3678) - 0.1627;
    return x * (0.2202 * x + 0.06448);
}

typedef struct { double z; int idx; } zpair;
static int zpair_cmp(const void* a, const void* b) {
    const zpair* p = (const zpair*)a; const zpair* q = (const zpair*)b;
    if (p->z < q->z) return -1;
    if (q->z < p->z) return 1;
    return (p->idx > q->idx) - (p->idx < q->idx);
}
static int int_cmp(const void* a, const void* b) { return (*(const int*)a > *(const int*)b) - (*(const int*)a < *(const int*)b); }

/* PolarCodes.cpp:45-56 (recursion unrolled bottom-up) + :59-74: keep the k LARGEST z. */
void po_inf_indexes_awgn(int n, int k, double sigma, int* out) {
    double* z = (double*)malloc(sizeof(double) * n);
    double* t = (double*)malloc(sizeof(double) * n);
    zpair* p = (zpair*)malloc(sizeof(zpair) * n);
    z[0] = 2. / (sigma * sigma);
    for (int s = 1; s < n; s *= 2) {
        for (int i = 0; i < s; i++) { t[2 * i] = awgn_phi(z[i]); t[2 * i + 1] = 2. * z[i]; }
        memcpy(z, t, sizeof(double) * 2 * s);
    }
    for (int i = 0; i < n; i++) { p[i].z = z[i]; p[i].idx = i; }
    qsort(p, n, sizeof(zpair), zpair_cmp);          /* std::sort on (z,index) pairs, :67 */
    for (int i = n - 1; i >= n - k; i--) out[n - i - 1] = p[i].idx;
    qsort(out, k, sizeof(int), int_cmp);
    free(z); free(t); free(p);
}

/* PolarCodes.cpp:18-29 + :77-92: Bhattacharyya recursion from 0.5, keep the k SMALLEST. */
void po_inf_indexes_bec(int n, int k, int* out) {
    double* z = (double*)malloc(sizeof(double) * n);
    double* t = (double*)malloc(sizeof(double) * n);
    zpair* p = (zpair*)malloc(sizeof(zpair) * n);
    z[0] = 0.5;
    for (int s = 1; s < n; s *= 2) {
        for (int i = 0; i < s; i++) { t[2 * i] = 2. * z[i] - z[i] * z[i]; t[2 * i + 1] = z[i] * z[i]; }
        memcpy(z, t, sizeof(double) * 2 * s);
    }
    for (int i = 0; i < n; i++) { p[i].z = z[i]; p[i].idx = i; }
    qsort(p, n, sizeof(zpair), zpair_cmp);
    for (int i = 0; i < k; i++) out[i] = p[i].idx;
    qsort(out, k, sizeof(int), int_cmp);
    free(z); free(t); free(p);
}

/* Encoder.cpp:43-55 with the all-zero frozen vector of PolarCodes.cpp:127: -1 info, 0 frozen. */
void po_frozen_mask(int n, const int* inf, int ninf, int* mask) {
    int u = 0;
    for (int i = 0; i < n; i++) {
        if (u < ninf && inf[u] == i) { mask[i] = -1; u++; } else mask[i] = 0;
    }
}

/* Encoder.cpp:17-33 (frozen values are all zero, PolarCodes.cpp:127). */
void po_add_frozen_bits(const int* xinf, int* x, int n, const int* inf, int ninf) {
    int u = 0;
    for (int i = 0; i < n; i++) {
        if (u < ninf && inf[u] == i) x[i] = xinf[u++]; else x[i] = 0;
    }
}

/* ------------------------------------------------------------------ encoder --------- */

/* Encoder.cpp:57-83.  The reference recursion [E(first)^E(second), E(second)] on halves is
 * the natural-order transform x = u F^{(x)m}; with withShuffle each node first de-interleaves
 * even/odd (Resources.cpp:148-158), which composes to a bit-reversal of the input index, i.e.
 * x = bitrev(u) F^{(x)m} (Arikan's B_N F^{(x)m}). */
void po_encode(int n, const int* u, int* x, int with_shuffle) {
    int m = 0;
    while ((1 << m) < n) m++;
    if (with_shuffle) {
        for (int i = 0; i < n; i++) {
            int r = 0;
            for (int b = 0; b < m; b++) if (i & (1 << b)) r |= 1 << (m - 1 - b);
            x[i] = u[r];
        }
    } else {
        for (int i = 0; i < n; i++) x[i] = u[i];
    }
    for (int s = 1; s < n; s *= 2)
        for (int i = 0; i < n; i++)
            if (!(i & s)) x[i] ^= x[i + s];
}

/* ------------------------------------------------------------------ CRC ------------- */

/* Resources.cpp:29-62 is the table of poly 0x31; :108-122 consumes one int per bit, 8 bits
 * MSB-first per step, init 0xFF.  Reads data[i+u] up to ceil8(len)-1 like the reference. */
unsigned po_crc8(const int* bits, int len) {
    unsigned crc = 0xFF;
    for (int i = 0; i < len; i += 8) {
        unsigned byte = 0;
        for (int u = 0; u < 8; u++) byte = ((byte << 1) ^ (unsigned)bits[i + u]) & 0xFF;
        crc ^= byte;
        for (int b = 0; b < 8; b++) crc = (crc & 0x80) ? ((crc << 1) ^ 0x31) & 0xFF : (crc << 1) & 0xFF;
    }
    return crc;
}

/* Resources.cpp:73-106 table of poly 0x1021; :124-138 init 0xFFFF, crc=(crc<<8)^T[(crc>>8)^byte]. */
unsigned po_crc16(const int* bits, int len) {
    unsigned crc = 0xFFFF;
    for (int i = 0; i < len; i += 8) {
        unsigned byte = 0;
        for (int u = 0; u < 8; u++) byte = ((byte << 1) ^ (unsigned)bits[i + u]) & 0xFF;
        crc ^= byte << 8;
        for (int b = 0; b < 8; b++) crc = (crc & 0x8000) ? ((crc << 1) ^ 0x1021) & 0xFFFF : (crc << 1) & 0xFFFF;
    }
    return crc;
}

/* PolarCodes.cpp:136-155: CRC appended MSB first at xinf[k..k+crc). */
void po_append_crc(int* xinf, int k, int crc) {
    unsigned r;
    if (crc == 8) r = po_crc8(xinf, k);
    else if (crc == 16) r = po_crc16(xinf, k);
    else return;
    for (int u = 1; u <= crc; u++) { xinf[k + crc - u] = (int)(r & 1u); r >>= 1; }
}

/* ------------------------------------------------------------------ SC min-sum ------ */

/* SCDecoderMinsum.cpp:14-16 — sign from a real product, so an underflowing product is "+". */
static double ms_f(double a, double b) {
    double aa = fabs(a), ab = fabs(b);
    double mn = (ab < aa) ? ab : aa;            /* std::min(aa, ab) */
    return mn * (a * b < 0 ? -1. : 1.);
}
/* SCDecoderMinsum.cpp:24-26 */
static double ms_g(double a, double b, int bit) { return b + a * (1 - bit * 2); }

typedef double (*f_fn)(double, double);

/* SCDecoder.cpp:19-25 */
static double ex_f(double y1, double y2) {
    double a = 2. * atanh(tanh(y1 / 2) * tanh(y2 / 2));
    if (isinf(a) || isnan(a)) {
        double aa = fabs(y1), ab = fabs(y2);
        double mn = (ab < aa) ? ab : aa;
        return mn * (a < 0 ? -1 : 1);
    }
    return a;
}
/* SCDecoder.cpp:27-36 */
static double ex_g(double y1, double y2, int b) { return b == 0 ? y1 + y2 : y2 - y1; }
typedef double (*g_fn)(double, double, int);

/* SCDecoderMinsum.cpp:32-66 / SCDecoder.cpp:38-72: adjacent-pair recursion, leaf at node
 * size 2, decision `> 0 ? 0 : 1`, interleaved partial sums.  llr/bits are per-depth scratch. */
static void sc_node(int size, int depth, double** llr, int** bits, const int* mask, int* sol, int* pos,
                    f_fn F, g_fn G) {
    double* in = llr[depth];
    int* out = bits[depth];
    if (size == 2) {
        int u1 = mask[*pos] != -1 ? mask[*pos] : (F(in[0], in[1]) > 0 ? 0 : 1);
        sol[(*pos)++] = u1;
        int u2 = mask[*pos] != -1 ? mask[*pos] : (G(in[0], in[1], u1) > 0 ? 0 : 1);
        sol[(*pos)++] = u2;
        out[0] = u1 ^ u2;
        out[1] = u2;
        return;
    }
    double* ch = llr[depth + 1];
    int* cb = bits[depth + 1];
    for (int i = 0; i < size; i += 2) ch[i / 2] = F(in[i], in[i + 1]);
    sc_node(size / 2, depth + 1, llr, bits, mask, sol, pos, F, G);
    for (int i = 0; i < size; i += 2) { out[i] = cb[i / 2]; ch[i / 2] = G(in[i], in[i + 1], out[i]); }
    sc_node(size / 2, depth + 1, llr, bits, mask, sol, pos, F, G);
    for (int i = 0; i < size; i += 2) { out[i] ^= cb[i / 2]; out[i + 1] = cb[i / 2]; }
}

static void sc_run(int n, const double* in0, const int* mask, int* sol, f_fn F, g_fn G) {
    int m = 0;
    while ((1 << m) < n) m++;
    double** llr = (double**)malloc(sizeof(double*) * (m + 1));
    int** bits = (int**)malloc(sizeof(int*) * (m + 1));
    for (int d = 0; d <= m; d++) {
        int s = (n >> d) > 2 ? (n >> d) : 2;
        llr[d] = (double*)malloc(sizeof(double) * s);
        bits[d] = (int*)malloc(sizeof(int) * s);
    }
    memcpy(llr[0], in0, sizeof(double) * n);
    int pos = 0;
    sc_node(n, 0, llr, bits, mask, sol, &pos, F, G);
    for (int d = 0; d <= m; d++) { free(llr[d]); free(bits[d]); }
    free(llr); free(bits);
}

/* SCDecoderMinsum.cpp:68-76 — raw channel values are used as LLRs (no scaling). */
void po_scms_decode(int n, const double* y, const int* mask, int* sol) { sc_run(n, y, mask, sol, ms_f, ms_g); }

/* SCDecoder.cpp:74-85 — LLR = 2y/sigma^2 (:15-17), inf clamped to DBL_MAX (:77-79). */
void po_sc_decode(int n, const double* y, const int* mask, double sigma, int* sol) {
    double* l = (double*)malloc(sizeof(double) * n);
    for (int i = 0; i < n; i++) {
        l[i] = y[i] * 2. / (sigma * sigma);
        if (isinf(l[i])) l[i] = DBL_MAX;
    }
    sc_run(n, l, mask, sol, ex_f, ex_g);
    free(l);
}

/* ------------------------------------------------------------------ SCL min-sum ----- */

/* SCLDecoderMinsum.cpp:40-42, :51-53 — fp32 */
static float scl_f(float a, float b) {
    float aa = fabsf(a), ab = fabsf(b);
    float mn = (ab < aa) ? ab : aa;
    return mn * (a * b < 0 ? -1.f : 1.f);
}
static float scl_g(float a, float b, int c) { return b + a * (float)(1 - c * 2); }
/* SCLDecoderMinsum.cpp:308-313 */
static float scl_pen(float llr, int u) {
    if ((u == 0 && llr < 0) || (u == 1 && llr > 0)) return fabsf(llr);
    return 0;
}

typedef struct {
    int n, m, L;
    float* P;   /* [L][2n]: layer l of slot p at P + p*2n + off[l] */
    int* C;     /* same layout */
    int* off;   /* [m+1] */
    float* met;
    int* active;
    int* stack; int sp;
} scl_t;

static float* SP(scl_t* s, int p, int l) { return s->P + (size_t)p * 2 * s->n + s->off[l]; }
static int* SC_(scl_t* s, int p, int l) { return s->C + (size_t)p * 2 * s->n + s->off[l]; }

/* clonePath :145-156 as an eager copy; the new slot is popped from the END of the free stack. */
static int scl_clone(scl_t* s, int p) {
    int q = s->stack[--s->sp];
    s->active[q] = 1;
    memcpy(s->P + (size_t)q * 2 * s->n, s->P + (size_t)p * 2 * s->n, sizeof(float) * 2 * s->n);
    memcpy(s->C + (size_t)q * 2 * s->n, s->C + (size_t)p * 2 * s->n, sizeof(int) * 2 * s->n);
    s->met[q] = s->met[p];
    return q;
}
/* killPath :157-168 */
static void scl_kill(scl_t* s, int p) { s->active[p] = 0; s->stack[s->sp++] = p; }

static int ctz_(int v) { int c = 0; while (!(v & 1)) { v >>= 1; c++; } return c; }

/* bit write (getArrayPointerCW :202-223) + iterativelyUpdateC :278-306 */
static void scl_setbit(scl_t* s, int p, int i, int u) {
    int m = s->m;
    if ((i & 1) == 0) { SC_(s, p, m)[0] = u; return; }
    int d = ctz_(i + 1);
    if (d > m) d = m;
    int* dest = SC_(s, p, m - d);
    int size = 1 << d;
    dest[size - 1] = u;
    int w = 1;                                   /* width of the finished right part */
    for (int l = m; l > m - d; l--) {
        int* c = SC_(s, p, l);                   /* finished left sibling of width w */
        for (int j = 0; j < w; j++) dest[size - 2 * w + j] = c[j] ^ dest[size - w + j];
        w *= 2;
    }
}

/* iterativelyCalcP :225-276 — half-split pairing (j, j+s). */
static void scl_calc_p(scl_t* s, int p, int i) {
    int m = s->m, n = s->n;
    int d = (i == 0) ? m - 1 : ctz_(i);
    if (d > m - 1) d = m - 1;
    int l = m - d;
    int sz = n >> l;
    if ((i >> d) & 1) {
        float* in = SP(s, p, l - 1); float* out = SP(s, p, l); int* c = SC_(s, p, l);
        for (int j = 0; j < sz; j++) out[j] = scl_g(in[j], in[j + sz], c[j]);
        l++; sz >>= 1;
    }
    for (; l <= m; l++, sz >>= 1) {
        float* in = SP(s, p, l - 1); float* out = SP(s, p, l);
        for (int j = 0; j < sz; j++) out[j] = scl_f(in[j], in[j + sz]);
    }
}

static int float_cmp(const void* a, const void* b) {
    float x = *(const float*)a, y = *(const float*)b;
    return (x > y) - (x < y);
}

/* continuePaths_UnfrozenBit :315-400 */
static void scl_fork(scl_t* s, int i, float* pf, float* sorted, int* keep) {
    int L = s->L, m = s->m, nact = 0;
    for (int p = 0; p < L; p++) {
        if (!s->active[p]) continue;
        float llr = SP(s, p, m)[0];
        pf[p] = s->met[p] + scl_pen(llr, 0);
        pf[p + L] = s->met[p] + scl_pen(llr, 1);
        nact++;
    }
    if (nact < L) {                               /* growth :329-341 (live `active` test) */
        for (int p = 0; p < L; p++) {
            if (!s->active[p]) continue;
            int q = scl_clone(s, p);
            scl_setbit(s, p, i, 0);
            scl_setbit(s, q, i, 1);
            s->met[p] = pf[p];
            s->met[q] = pf[p + L];
        }
        return;
    }
    for (int p = 0; p < L; p++) { sorted[2 * p] = pf[p]; sorted[2 * p + 1] = pf[p + L]; }
    qsort(sorted, 2 * L, sizeof(float), float_cmp);
    float med = sorted[L];                        /* nth_element(...)[L], :343-344 */
    int c = 0;
    for (int p = 0; p < L; p++) {                 /* pass 1 :346-361 */
        keep[p] = (pf[p] < med && c < L); c += keep[p];
        keep[p + L] = (pf[p + L] < med && c < L); c += keep[p + L];
    }
    for (int p = 0; p < L && c < L; p++) {        /* pass 2 :362-371 */
        if (pf[p] == med) { keep[p] = 1; c++; }
        if (pf[p + L] == med && c < L) { keep[p + L] = 1; c++; }
    }
    for (int p = 0; p < L; p++) if (!keep[p] && !keep[p + L]) scl_kill(s, p);   /* :372-376 */
    for (int p = 0; p < L; p++) {                 /* :377-398 */
        if (!s->active[p] || (!keep[p] && !keep[p + L])) continue;
        if (keep[p] && keep[p + L]) {
            int q = scl_clone(s, p);
            scl_setbit(s, p, i, 0);
            scl_setbit(s, q, i, 1);
            s->met[p] = pf[p];
            s->met[q] = pf[p + L];
        } else if (keep[p]) {
            scl_setbit(s, p, i, 0); s->met[p] = pf[p];
        } else {
            scl_setbit(s, p, i, 1); s->met[p] = pf[p + L];
        }
    }
}

/* SCLDecoderMinsum.cpp:401-477 */
int po_scl_decode(int n, int L, int k, int crc, const double* y, const int* mask, int* sol) {
    scl_t s;
    int m = 0;
    while ((1 << m) < n) m++;
    if (L < 1 || (L & (L - 1))) return -1;
    s.n = n; s.m = m; s.L = L;
    s.P = (float*)calloc((size_t)L * 2 * n, sizeof(float));
    s.C = (int*)calloc((size_t)L * 2 * n, sizeof(int));
    s.off = (int*)malloc(sizeof(int) * (m + 1));
    s.met = (float*)calloc(L, sizeof(float));
    s.active = (int*)calloc(L, sizeof(int));
    s.stack = (int*)malloc(sizeof(int) * L);
    float* pf = (float*)calloc(2 * L, sizeof(float));
    float* sorted = (float*)calloc(2 * L, sizeof(float));
    int* keep = (int*)calloc(2 * L, sizeof(int));
    int* uhat = (int*)malloc(sizeof(int) * n);
    if (!s.P || !s.C || !uhat) return -1;
    int o = 0;
    for (int l = 0; l <= m; l++) { s.off[l] = o; o += n >> l; }
    for (int p = 0; p < L; p++) s.stack[p] = p;   /* free stack [0..L-1], :93-96 */
    s.sp = L;
    int p0 = s.stack[--s.sp];                     /* assignInitialPath :133-144 -> slot L-1 */
    s.active[p0] = 1; s.met[p0] = 0;
    for (int j = 0; j < n; j++) SP(&s, p0, 0)[j] = (float)y[j];   /* :406-408 */
    for (int i = 0; i < n; i++) {
        for (int p = 0; p < L; p++) {
            if (!s.active[p]) continue;
            scl_calc_p(&s, p, i);
            if (mask[i] != -1) {                  /* :413-418 */
                scl_setbit(&s, p, i, mask[i]);
                s.met[p] += scl_pen(SP(&s, p, m)[0], mask[i]);
            }
        }
        if (mask[i] == -1) scl_fork(&s, i, pf, sorted, keep);
    }
    /* :432-473 — slots in index order, first CRC pass wins, else the last slot's word stays */
    for (int p = 0; p < L; p++) {
        po_encode(n, SC_(&s, p, 0), uhat, 0);
        int j = 0;
        for (int u = 0; u < n; u++) if (mask[u] == -1) sol[j++] = uhat[u];
        int ok = 0;
        if (crc == 8 || crc == 16) {
            unsigned r = crc == 8 ? po_crc8(sol, k) : po_crc16(sol, k);
            ok = 1;
            for (int u = 1; u <= crc; u++) {
                if (sol[k + crc - u] != (int)(r & 1u)) { ok = 0; break; }
                r >>= 1;
            }
        }
        if (ok) break;
    }
    free(s.P); free(s.C); free(s.off); free(s.met); free(s.active); free(s.stack);
    free(pf); free(sorted); free(keep); free(uhat);
    return 0;
}

/* ------------------------------------------------------------------ 8-bit min-sum SC ---
 * The reference sketches an 8-bit decoder and never wires it up: Decodef_char / Decodeg_char
 * (SCDecoderMinsum.cpp:17-30) and the quantiser scale `fraction = 2 << qFraction`
 * (SCLDecoderMinsum.cpp:405, conf.txt:9).  This restatement is the definition of the product's
 * PCB_PREC_INT8 mode (a mode of its own: NOT bit-exact with the double decoder, reported with its
 * own FER curve): the tree walk of SCDecoderMinsum.cpp:32-66 on int8 LLRs
 *   q(y)     = clamp(rint(y * (2 << qFraction)), -127, 127)          (round half to even)
 *   f(a, b)  = Decodef_char(a, b): min(|a|, |b|), negative iff the sign bits differ
 *   g(a,b,c) = clamp(b + a * (1 - 2c), -127, 127)   (Decodeg_char is the same sum WITHOUT the clamp: it
 *              wraps in 8 bits, which no fixed-point decoder can live with; the clamp is this mode's
 *              one documented deviation from the sketch)
 *   decision = llr > 0 ? 0 : 1. */
int po_q8(double y, int qfraction) {
    double v = rint(y * (double)(2 << qfraction));
    if (!(v >= -127.)) v = -127.;      /* (NaN -> -127) */
    if (v > 127.) v = 127.;
    return (int)v;
}
int po_f8(int r1, int r2) {
    int m1 = r1 > 0 ? r1 : -r1, m2 = r2 > 0 ? r2 : -r2;
    int mm = m1 < m2 ? m1 : m2;
    return (((r1 ^ r2) & 0x80) == 0) ? mm : -mm;
}
int po_g8(int r1, int r2, int b) {
    int v = r2 + r1 * (1 - b * 2);
    return v > 127 ? 127 : v < -127 ? -127 : v;
}
static void sc8_node(int size, int depth, int** llr, int** bits, const int* mask, int* sol, int* pos) {
    int* in = llr[depth];
    int* out = bits[depth];
    if (size == 2) {
        int u1 = mask[*pos] != -1 ? mask[*pos] : (po_f8(in[0], in[1]) > 0 ? 0 : 1);
        sol[(*pos)++] = u1;
        int u2 = mask[*pos] != -1 ? mask[*pos] : (po_g8(in[0], in[1], u1) > 0 ? 0 : 1);
        sol[(*pos)++] = u2;
        out[0] = u1 ^ u2;
        out[1] = u2;
        return;
    }
    int* ch = llr[depth + 1];
    int* cb = bits[depth + 1];
    for (int i = 0; i < size; i += 2) ch[i / 2] = po_f8(in[i], in[i + 1]);
    sc8_node(size / 2, depth + 1, llr, bits, mask, sol, pos);
    for (int i = 0; i < size; i += 2) { out[i] = cb[i / 2]; ch[i / 2] = po_g8(in[i], in[i + 1], out[i]); }
    sc8_node(size / 2, depth + 1, llr, bits, mask, sol, pos);
    for (int i = 0; i < size; i += 2) { out[i] ^= cb[i / 2]; out[i + 1] = cb[i / 2]; }
}
int po_sc8_decode_batch(int n, int qfraction, const double* y, const int* mask, int* sol, int frames) {
    int m = 0;
    while ((1 << m) < n) m++;
    int** llr = (int**)malloc(sizeof(int*) * (m + 1));
    int** bits = (int**)malloc(sizeof(int*) * (m + 1));
    for (int d = 0; d <= m; d++) {
        int s = (n >> d) > 2 ? (n >> d) : 2;
        llr[d] = (int*)malloc(sizeof(int) * s);
        bits[d] = (int*)malloc(sizeof(int) * s);
    }
    for (int f = 0; f < frames; f++) {
        for (int i = 0; i < n; i++) llr[0][i] = po_q8(y[(size_t)f * n + i], qfraction);
        int pos = 0;
        sc8_node(n, 0, llr, bits, mask, sol + (size_t)f * n, &pos);
    }
    for (int d = 0; d <= m; d++) { free(llr[d]); free(bits[d]); }
    free(llr); free(bits);
    return 0;
}

/* ------------------------------------------------------------------ batch helpers --- */

int po_decode_batch(int decoder, int n, int L, int k, int crc, double sigma, const double* y,
                    const int* mask, int* sol, int frames) {
    int stride = decoder == 3 ? k + crc : n;
    int* tmp = (int*)malloc(sizeof(int) * n);
    for (int f = 0; f < frames; f++) {
        const double* c = y + (size_t)f * n;
        if (decoder == 1) po_sc_decode(n, c, mask, sigma, tmp);
        else if (decoder == 2) po_scms_decode(n, c, mask, tmp);
        else if (po_scl_decode(n, L, k, crc, c, mask, tmp)) { free(tmp); return -1; }
        memcpy(sol + (size_t)f * stride, tmp, sizeof(int) * stride);
    }
    free(tmp);
    return 0;
}

double po_time_decode(int decoder, int n, int L, int k, int crc, double sigma, const double* y,
                      const int* mask, int frames, int reps) {
    int stride = decoder == 3 ? k + crc : n;
    int* sol = (int*)malloc(sizeof(int) * (size_t)stride * frames);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < reps; r++) po_decode_batch(decoder, n, L, k, crc, sigma, y, mask, sol, frames);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(sol);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
