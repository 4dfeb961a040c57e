// Host-side code construction, CRC and frame assembly — runs once per SNR point, tiny.
// Mirrors the observable results of (paths relative to /root/reference/PolarCodes/):
//   sigma                 PolarCodes.cpp:123
//   InfIndexesAWGN/BEC    PolarCodes.cpp:18-92   (reliability recursion + selection, ties by index)
//   FrozenBits            Encoder.cpp:43-55
//   AddFrozenBits         Encoder.cpp:17-33
//   Crc8 / Crc16          Resources.cpp:108-138, append loop PolarCodes.cpp:136-155
// Must be compiled with -ffp-contract=off: a fused multiply-add in the AWGN polynomial can move
// near-tied reliabilities and change the frozen set.
#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

#include "../../include/polar_b200.h"
#include "errors.h"

namespace {

// 4-piece approximation of the check-node mean update (PolarCodes.cpp:31-44)
inline double awgn_check_update(double x) {
    if (x > 12) return 0.9861 * x - 2.3152;
    if (x > 3.5) return x * (0.009005 * x + 0.7694) - 0.9507;
    if (x > 1) return x * (0.062883 * x + 0.3678) - 0.1627;
    return x * (0.2202 * x + 0.06448);
}

bool pow2(int n) { return n >= 1 && (n & (n - 1)) == 0; }

// Expand z[0] level by level in place (children of i at 2i, 2i+1); walking i downwards keeps
// every parent readable until both children are written.
template <class Minus, class Plus>
std::vector<double> reliabilities(int n, double root, Minus minus, Plus plus) {
    std::vector<double> z((size_t)n);
    z[0] = root;
    for (int s = 1; s < n; s *= 2)
        for (int i = s - 1; i >= 0; --i) {
            const double v = z[i];
            z[2 * i + 1] = plus(v);
            z[2 * i] = minus(v);
        }
    return z;
}

// indices ordered by (z, index) ascending — what std::sort on pair<double,int> yields
std::vector<int> order_by_reliability(const std::vector<double>& z) {
    std::vector<int> idx(z.size());
    std::iota(idx.begin(), idx.end(), 0);
    std::sort(idx.begin(), idx.end(), [&](int a, int b) { return z[a] < z[b] || (!(z[b] < z[a]) && a < b); });
    return idx;
}

}  // namespace

extern "C" {

double pcb_sigma(int n, int k, double snr_db) { return std::sqrt(n / (2. * k * std::pow(10, snr_db / 10))); }

int pcb_inf_indexes_awgn(int n, int k, double sigma, int* out) {
    if (!pow2(n) || k < 0 || k > n || !out) return pcb::fail(PCB_EINVAL, "pcb_inf_indexes_awgn: bad n/k");
    auto z = reliabilities(n, 2. / (sigma * sigma), awgn_check_update, [](double v) { return 2. * v; });
    auto idx = order_by_reliability(z);
    std::vector<int> pick(idx.end() - k, idx.end());   // the k LARGEST
    std::sort(pick.begin(), pick.end());
    std::copy(pick.begin(), pick.end(), out);
    return PCB_OK;
}

int pcb_inf_indexes_bec(int n, int k, int* out) {
    if (!pow2(n) || k < 0 || k > n || !out) return pcb::fail(PCB_EINVAL, "pcb_inf_indexes_bec: bad n/k");
    auto z = reliabilities(
        n, 0.5, [](double v) { return 2. * v - v * v; }, [](double v) { return v * v; });
    auto idx = order_by_reliability(z);
    std::vector<int> pick(idx.begin(), idx.begin() + k);   // the k SMALLEST Bhattacharyya parameters
    std::sort(pick.begin(), pick.end());
    std::copy(pick.begin(), pick.end(), out);
    return PCB_OK;
}

int pcb_frozen_mask(int n, const int* inf, int ninf, int* mask) {
    if (n < 1 || !mask || (ninf && !inf)) return pcb::fail(PCB_EINVAL, "pcb_frozen_mask: bad arguments");
    std::fill(mask, mask + n, 0);
    for (int t = 0; t < ninf; t++) {
        if (inf[t] < 0 || inf[t] >= n) return pcb::fail(PCB_EINVAL, "pcb_frozen_mask: index out of range");
        mask[inf[t]] = -1;
    }
    return PCB_OK;
}

int pcb_add_frozen_bits(const int* xinf, int* x, int n, const int* inf, int ninf) {
    if (n < 1 || !x || !xinf || !inf) return pcb::fail(PCB_EINVAL, "pcb_add_frozen_bits: bad arguments");
    std::fill(x, x + n, 0);
    for (int t = 0; t < ninf; t++) x[inf[t]] = xinf[t];
    return PCB_OK;
}

// bitwise LFSR, one int per bit, bytes assembled MSB first (Resources.cpp:108-122)
unsigned pcb_crc8(const int* bits, int len) {
    unsigned r = 0xFF;
    for (int i = 0; i < len; i++) {
        r ^= (unsigned)(bits[i] & 1) << 7;
        r = (r & 0x80) ? ((r << 1) ^ 0x31) & 0xFF : (r << 1) & 0xFF;
    }
    return r;
}

// CCITT 0x1021, init 0xFFFF (Resources.cpp:124-138)
unsigned pcb_crc16(const int* bits, int len) {
    unsigned r = 0xFFFF;
    for (int i = 0; i < len; i++) {
        r ^= (unsigned)(bits[i] & 1) << 15;
        r = (r & 0x8000) ? ((r << 1) ^ 0x1021) & 0xFFFF : (r << 1) & 0xFFFF;
    }
    return r;
}

int pcb_append_crc(int* xinf, int k, int crc) {
    if (crc == 0) return PCB_OK;
    if ((crc != 8 && crc != 16) || k % 8 || !xinf) return pcb::fail(PCB_EINVAL, "pcb_append_crc: crc in {0,8,16}, k % 8 == 0");
    const unsigned r = crc == 8 ? pcb_crc8(xinf, k) : pcb_crc16(xinf, k);
    for (int t = 0; t < crc; t++) xinf[k + t] = (int)((r >> (crc - 1 - t)) & 1u);
    return PCB_OK;
}


// Algorithmic ALU work of one frame under the SURVEY.md section 8(d) convention, counted exactly from the
// frozen mask: f = 4 lane-ops, g = 2, path-metric update 2 per live path and phase, one compare-exchange
// of the (bitonic-convention) median select 3, one per packed 32-bit word of partial-sum XORs.  The list
// is a_i = min(L, 2^(information positions before i)) paths wide at phase i; phase i recomputes layers
// m-d..m with d = min(m-1, ctz i) (d = m-1 at i = 0), the first one a g-layer iff bit d of i is set.
int pcb_lane_ops_per_frame(const int* frozen_mask, int n, int L, int decoder, double* lane_ops, double* llr_updates) {
    const bool list = decoder == PCB_DEC_SCLMINSUM;     // the SC decoders keep no metric and select nothing
    if (!list) L = 1;
    if (!frozen_mask || !pow2(n) || n < 2 || L < 1 || !lane_ops) return pcb::fail(PCB_EINVAL, "pcb_lane_ops_per_frame: bad argument");
    int m = 0;
    while ((1 << m) < n) m++;
    double f = 0, g = 0, sa = 0;
    int seen = 0, ninfo = 0;
    for (int i = 0; i < n; i++) {
        const double a = seen >= 30 ? L : std::min<double>(L, (double)(1u << seen));
        const int d = i == 0 ? m - 1 : std::min(m - 1, __builtin_ctz((unsigned)i));
        const double size = (double)(1u << d);
        if ((i >> d) & 1) { g += a * size; f += a * (size - 1); }
        else f += a * (2 * size - 1);
        sa += a;
        if (frozen_mask[i] == -1) { seen++; ninfo++; }
    }
    int l2 = 0;
    while ((1 << l2) < 2 * L) l2++;
    const double ce = list ? (double)ninfo * L * l2 * (l2 + 1) / 2 : 0;
    const double x = std::floor((f + g) / 2 / 32);
    *lane_ops = 4 * f + 2 * g + (list ? 2 * sa : 0) + 3 * ce + x;
    if (llr_updates) *llr_updates = f + g;
    return PCB_OK;
}

}  // extern "C"
