// C ABI (include/polar_b200.h): handle, code upload, batched encode / decode / Monte-Carlo.
// There is no CPU fallback anywhere in this file: every compute entry point launches CUDA
// kernels and fails loudly if the device or the launch is unavailable.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>
#include <immintrin.h>
#include <sched.h>

#include "../../include/polar_b200.h"
#include "aux_kernels.h"
#include "decode_launch.h"
#include "errors.h"
#include "nccl_dl.h"

namespace pcb {
static thread_local char g_err[512] = "";
int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
cudaError_t decode_generic_occupancy(int L, int mode, int fp64, int wpb, size_t smem, int* bps);
// min-sum SC kernel with 32 frames per warp (decode_sc32.cu)
bool sc32_supported(int m);
cudaError_t sc32_occupancy(int m, int fp64, int* bps, int* bytes);
size_t sc32_yt_bytes(int n, int B, int fp64);
size_t sc32_scratch_bytes(int blocks, int fp64);
bool sc32_is_big(int m);
size_t sc32_big_scratch_bytes(int n, int blocks, int fp64);
cudaError_t launch_decode_sc32(const DecodeArgs& a, const uint32_t* info_w, const uint32_t* fv_w, void* yt, void* scratch,
                               uint32_t* tmp_bits, int fp64, int exact, int blocks, int sm_count, cudaStream_t st);
cudaError_t launch_sc_expand(const uint32_t* bits, int* sol, size_t total, int sm_count, cudaStream_t st);
// SC decoders, one tile of 32 frames per block with the whole level store on chip and a pruned tree walk (decode_sct.cu)
bool sct_supported(int m);
cudaError_t sct_occupancy(int m, int fp64, int exact, int W, int vs, int* bps, size_t* smem, size_t* glob);
int sct_build_ops(int n, int m, int W, const uint32_t* info_w, const uint32_t* fv_w, uint32_t* ops_out, int cap, uint32_t* pops_out, int pcap,
                  int* npops);
cudaError_t launch_decode_sct(const DecodeArgs& a, const uint32_t* d_ops, int nops, const uint32_t* d_pops, int npops, uint32_t* bits, void* glev,
                              int fp64, int exact, int W, int vs, int blocks, cudaStream_t st);
// 8-bit min-sum SC, four LLRs per register (decode_sc8.cu; PCB_PREC_INT8)
bool sc8_supported(int m);
size_t sc8_yq_bytes(int n, int B);
cudaError_t sc8_occupancy(int m, int* bps, size_t* smem, int* vs);
int sc8_build_ops(int n, int m, const uint32_t* info_w, const uint32_t* fv_w, uint32_t* ops_out, int cap);
size_t sc8_glev_bytes(int n, int blocks, int vs);
cudaError_t launch_decode_sc8(const DecodeArgs& a, int qfraction, const uint32_t* d_ops, int nops, uint32_t* yq, uint32_t* glev, uint32_t* bits, int blocks,
                              int sm_count, int vs, cudaStream_t st);
// lane-per-path list kernel, 32 / L frames per warp (decode_scl8q.cu): L = 8 and L = 32
bool scl8q_supported(int m, int L);
int scl8q_frames_per_warp(int L);
int scl8q_warps_per_block(int m, int L);
cudaError_t launch_decode_scl8q(const DecodeArgs& a, int L, const uint32_t* info_w, uint32_t* xhat, void* l3_scratch, int* next_quad,
                                const double* y64, int nf, float* yscr, int check_range, const Scl8qGen* gen, int blocks, cudaStream_t st);
size_t scl8q_scratch_per_block(int m, int L);
cudaError_t scl8q_occupancy(int m, int L, int* bps, int* warp_bytes);
}  // namespace pcb

using namespace pcb;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(PCB_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// Host side of pcb_decode_batch for the float decoders: the reference narrows every channel value
// to float before it decodes (SCLDecoderMinsum.cpp:406-408), so the narrowing is done on the host
// cores, chunk by chunk into pinned staging buffers, and only 4 bytes per value cross PCIe (the
// link, not the GPU, bounds this call).  Same rounding as the device cast (round to nearest even,
// denormals kept).  This is data movement of the product path, not a decoder: no CPU fallback.
struct HostNarrow {
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv, done_cv;
    const double* src = nullptr;
    float* dst = nullptr;
    size_t cnt = 0;
    int gen = 0, pending = 0, T = 0;
    bool stop = false;

    static int usable_cpus() {     // the CPUs this process may run on (a container sees fewer than the host has)
        cpu_set_t set;
        CPU_ZERO(&set);
        if (sched_getaffinity(0, sizeof(set), &set) == 0 && CPU_COUNT(&set) > 0) return CPU_COUNT(&set);
        return (int)std::thread::hardware_concurrency();
    }
    __attribute__((target("avx"))) static size_t narrow_avx(const double* s, float* d, size_t n) {   // d 32-byte aligned
        size_t i = 0;
        for (; i + 8 <= n; i += 8) {
            const __m128 lo = _mm256_cvtpd_ps(_mm256_loadu_pd(s + i)), hi = _mm256_cvtpd_ps(_mm256_loadu_pd(s + i + 4));
            _mm256_stream_ps(d + i, _mm256_insertf128_ps(_mm256_castps128_ps256(lo), hi, 1));
        }
        return i;
    }
    static void narrow(const double* s, float* d, size_t n) {   // d 32-byte aligned; the host never reads the staging buffer
        size_t i = 0;
        if (__builtin_cpu_supports("avx")) i = narrow_avx(s, d, n);
        for (; i + 4 <= n; i += 4) {
            const __m128 lo = _mm_cvtpd_ps(_mm_loadu_pd(s + i)), hi = _mm_cvtpd_ps(_mm_loadu_pd(s + i + 2));
            _mm_stream_ps(d + i, _mm_movelh_ps(lo, hi));
        }
        for (; i < n; i++) d[i] = (float)s[i];
        _mm_sfence();
    }
    void worker(int id) {
        int seen = 0;
        for (;;) {
            std::unique_lock<std::mutex> lk(m);
            cv.wait(lk, [&] { return stop || gen != seen; });
            if (stop) return;
            seen = gen;
            const double* s = src; float* d = dst; const size_t n = cnt;
            lk.unlock();
            const size_t per = ((n + T - 1) / T + 15) & ~(size_t)15;
            const size_t a = std::min(n, (size_t)id * per), b = std::min(n, a + per);
            if (b > a) narrow(s + a, d + a, b - a);
            lk.lock();
            if (--pending == 0) done_cv.notify_one();
        }
    }
    void start() {
        if (T) return;
        const char* e = getenv("PCB_HOST_THREADS");
        int t;
        if (e && *e) {
            t = atoi(e);
        } else {
            // the ranks of one node (torchrun) share its CPUs; one CPU per rank stays with the calling
            // thread and the driver
            const char* lw = getenv("LOCAL_WORLD_SIZE");
            const int ranks = (lw && atoi(lw) > 0) ? atoi(lw) : 1;
            t = usable_cpus() / ranks - (ranks > 1 ? 1 : 2);
        }
        T = t < 1 ? 1 : t > 64 ? 64 : t;
        for (int i = 0; i < T; i++) th.emplace_back([this, i] { worker(i); });
    }
    void run(const double* s, float* d, size_t n) {
        std::unique_lock<std::mutex> lk(m);
        src = s; dst = d; cnt = n; pending = T; gen++;
        cv.notify_all();
        done_cv.wait(lk, [&] { return pending == 0; });
    }
    ~HostNarrow() {
        { std::lock_guard<std::mutex> lk(m); stop = true; }
        cv.notify_all();
        for (auto& t : th) t.join();
    }
};

struct pcb_handle {
    int device = 0, n = 0, m = 0, k = 0, crc = 0, L = 1, decoder = 0, precision = 0;
    int stride = 0, EW = 0, fp64 = 0, sizeofT = 4;
    int ninfo = 0;                  // information positions of the uploaded mask (decoders 1, 2: any count)
    bool code_set = false;
    double sigma = 1.0;
    int sm_count = 0, wpb = 4, bps = 1, ls = 1, warp_smem = 0;
    bool fast8 = false;             // tuned lane-per-path list kernel (decode_scl8q.cu) instead of the generic one
    // launch shape of the tuned kernel (the fields above always describe the generic kernel, which
    // also re-decodes the frames the tuned kernel hands over)
    int f_bps = 0, f_warp_smem = 0;
    bool quad = false;              // same as fast8 (kept: the launch code reads better with both names)
    bool sc32 = false;              // min-sum SC: lane-per-frame kernel (decode_sc32.cu) instead of the generic one
    int sc32_bps = 0, sc32_smem = 0;
    bool sc8 = false;               // decoder 2 in PCB_PREC_INT8: decode_sc8.cu
    int qfrac = 0, sc8_bps = 0, sc8_nops = 0;
    size_t sc8_smem = 0, yq_cap = 0;
    int sc8_vs = 2;
    uint32_t *d_sc8_ops = nullptr, *d_yq = nullptr, *d_sc8_glev = nullptr;
    bool sct = false;               // SC decoders: tile-per-block kernel (decode_sct.cu) instead of sc32
    int sct_bps = 0, sct_nops = 0, sct_npops = 0, sct_W = 4, sct_vs = 2;
    size_t sct_smem = 0, sct_glob = 0;
    uint32_t* d_sct_ops = nullptr;  // the pruned tree walk of the uploaded frozen pattern (pcb_set_code): warp 0's list, then the other warps'
    void* d_sct_glev = nullptr;     // levels 2 .. vs-1 (and the heap for n > 2048) of every resident block
    void* d_yt = nullptr;           // sc32: transposed channel values of the batch
    void* d_sc32_scratch = nullptr; // sc32: size-256 level of every resident block (stays in L2)
    uint32_t* d_sc32_bits = nullptr; // sc32: packed decisions when the caller wants one int per bit
    size_t sc32_bits_cap = 0;
    size_t yt_cap = 0;
    void* d_qscratch = nullptr;     // quad kernel: layer-3 arrays of every resident block (stays in L2)
    uint32_t* d_synd = nullptr;     // quad kernel: rows of the CRC syndrome matrix (16 x n/32 words), see pcb_set_code
    uint32_t synd_c0 = 0;
    uint32_t* d_gcrc = nullptr;     // quad kernel, Monte-Carlo mode: payload -> CRC map, [ceil((k+crc)/32)][16] words
    uint32_t gcrc_c0 = 0;
    uint32_t* d_xhat = nullptr;     // quad kernel: codeword estimates of every slot of the batch, the input of scl_select_kernel
    size_t xhat_cap = 0;            // (frames)
    uint32_t* d_skip = nullptr;     // Monte-Carlo mode: frames the select kernel must not count
    size_t skip_cap = 0;
    unsigned long long* d_mc_redo = nullptr;   // [0] count + ids of generated frames outside the tuned kernel's input range
    float* d_y32 = nullptr;         // PCB_SEPARATE_PREPASS only: double channel values converted once per batch
    float* d_yscr = nullptr;        // quad kernel: the float rows of every resident warp's frames (narrowed by the warp itself)
    size_t y32_cap = 0;
    int* d_redo = nullptr;          // [0] = count, [1] = work counter of the quad kernel, [2..] = frame list
    size_t redo_cap = 0;
    size_t scratch_per_warp = 0;
    cudaStream_t sc = nullptr, sin = nullptr, sout = nullptr;
    cudaEvent_t ev_in[2] = {}, ev_k[2] = {}, ev_out[2] = {}, t0 = nullptr, t1 = nullptr, ev_caller = nullptr;
    uint32_t *d_info_mask = nullptr, *d_frozen_val = nullptr;
    uint16_t* d_info_pos = nullptr;
    unsigned char* d_scratch = nullptr;
    unsigned char* d_noalloc = nullptr;   // pcb_decode_batch_dev_noalloc: device address of decode_generic.cu's file-scope scratch array
    size_t noalloc_bytes = 0;
    // host-buffer pipeline (two slots)
    size_t slot_frames = 0;
    void* d_y[2] = {};
    int* d_sol[2] = {};
    uint32_t* d_solb[2] = {};
    float* h_stage[2] = {};         // pinned staging of the narrowed channel values (host pipeline, float decoders)
    size_t stage_frames = 0;
    HostNarrow* narrow = nullptr;
    const double* mix_tail = nullptr;   // set by decode_host for one launch(): frames >= mix_nf of the batch are still
    int mix_nf = 0;                     // double, at this device address (the list-8 pre-pass converts them)
    double narrow_frac = 0.75;          // share of a chunk narrowed on the host (see decode_host)
    bool shared_host = false;           // other ranks of the job run on this host (LOCAL_WORLD_SIZE > 1)
    double pc_dir = 1.0, pc_prev = 0, pc_frames = 0, pc_time = 0;
    int pc_chunks = 0;
    double rate_cpu = 0, rate_h2d = 0;  // values/s narrowed by the host cores, bytes/s of the H2D copies
    cudaEvent_t ev_c0[2] = {}, ev_c1[2] = {};
    size_t copy_bytes[2] = {};
    // Monte-Carlo buffers
    size_t mc_frames = 0;
    float* d_ymc = nullptr;
    uint32_t *d_tx = nullptr, *d_dec = nullptr;
    unsigned long long* d_counters = nullptr;
    // encode buffers
    size_t enc_frames = 0;
    int *d_eu = nullptr, *d_ex = nullptr;
    size_t encp_frames = 0;
    uint32_t *d_epu[2] = {}, *d_epx[2] = {};   // packed encode: two slots
    uint64_t launches = 0;
    // counter reduction across the ranks of a box (pcb_comm_init): NCCL communicator + device buffer
    void* comm = nullptr;
    int comm_rank = 0, comm_world = 1;
    unsigned long long* d_red = nullptr;   // [8]: frames, frame errors, bit errors (+ caller's extras)
    uint64_t collectives = 0;
    std::vector<uint32_t> h_info_w, h_fv_w;   // host copies of the mask words (kernel parameters of the list-8 kernel)
};

// resident grid of the generic kernel; with the tuned kernel in front it only sees the rare
// out-of-range frames, so one block per SM is plenty (and keeps its scratch small)
static int max_blocks(const pcb_handle* h) { return h->sm_count * (h->fast8 ? 1 : h->bps); }
static int fast_blocks(const pcb_handle* h) { return h->sm_count * h->f_bps; }

static int ensure_device(pcb_handle* h) {
    CK(cudaSetDevice(h->device));
    return PCB_OK;
}

// choose the LLR-layer placement and the launch shape: the smallest first_smem_layer whose
// per-warp slice fits the budget, then the warps-per-block with the best occupancy.
static int configure(pcb_handle* h) {
    const char* wenv = getenv("PCB_WPB");
    h->fast8 = h->decoder == 3 && scl8q_supported(h->m, h->L) && !getenv("PCB_FORCE_GENERIC");
    h->quad = false;
    if (h->fast8) {
        // one lane per path, 32 / L frames per warp (decode_scl8q.cu)
        int bps = 0, wb = 0;
        if (scl8q_occupancy(h->m, h->L, &bps, &wb) != cudaSuccess || bps == 0) { cudaGetLastError(); h->fast8 = false; }
        else {
            const char* benv = getenv("PCB_Q_BPS");     // experiment: fewer resident warps per SM
            if (benv && atoi(benv) > 0 && atoi(benv) < bps) bps = atoi(benv);
            h->f_bps = bps; h->f_warp_smem = wb; h->quad = true;
        }
    }
    h->sc8 = h->decoder == 2 && h->precision == PCB_PREC_INT8;
    if (h->sc8) {
        if (!sc8_supported(h->m) || sc8_occupancy(h->m, &h->sc8_bps, &h->sc8_smem, &h->sc8_vs) != cudaSuccess || h->sc8_bps == 0) {
            cudaGetLastError();
            return fail(PCB_EINVAL, "PCB_PREC_INT8 needs n in [64, 4096]");
        }
        const char* be = getenv("PCB_8_BPS");
        if (be && atoi(be) > 0 && atoi(be) < h->sc8_bps) h->sc8_bps = atoi(be);
    }
    h->sc32 = !h->sc8 && (h->decoder == 2 || h->decoder == 1) && sc32_supported(h->m) && !getenv("PCB_FORCE_GENERIC");
    if (h->sc32 && (sc32_occupancy(h->m, h->fp64, &h->sc32_bps, &h->sc32_smem) != cudaSuccess || h->sc32_bps == 0)) {
        cudaGetLastError(); h->sc32 = false;
    }
    h->sct = !h->sc8 && (h->decoder == 2 || h->decoder == 1) && sct_supported(h->m) && !getenv("PCB_FORCE_GENERIC") && !getenv("PCB_SC_LANE");   // PCB_SC_LANE=1: the lane-per-frame kernel (A/B)
    // n > 2048: every level above the last 120 rows and the heap live in the block's global slice; four warps per tile and
    // 740 tiles in flight move 4.4 MB per frame where the lane-per-frame kernel moves 5.9 (config E: 1.17 vs 0.95 M frames/s)
    if (h->sct) {
        // Launch shape (tools/sct_sweep.py): warps per tile, and the first level kept in shared memory -- the rest of the
        // level store lives in the block's global slice (L2).  What decides is how many tiles are resident: the leaf walk
        // of a tile runs on one warp, the other tiles of the SM fill its gaps, so the shared-memory levels are kept to 16 KB.
        const bool exact = h->decoder == 1;
        const char* we = getenv("PCB_T_W");
        const int wv = we ? atoi(we) : 0;
        // measured best shapes (B200, tools/sct_sweep.py): {W, VS} per (arithmetic, n); elsewhere 16 KB of shared-memory levels
        int tw = exact ? (h->n <= 256 ? 1 : h->n <= 512 ? 2 : 4) : h->n <= 256 ? 1 : h->n == 512 ? 2 : 4, tvs = 0;
        if (exact) tvs = 4;
        else if (h->fp64) tvs = h->n == 2048 ? 5 : h->n <= 256 ? 3 : 4;
        else { if (h->n >= 1024) tvs = 4; else if (h->n <= 256) tvs = 2; }     // (fp32, n = 1024: one warp / VS 5 is 13 % faster at 2^18 frames and 10 % slower at 2^20)
        h->sct_W = (wv == 1 || wv == 2 || wv == 4 || wv == 8) ? wv : tw;
        if (h->n < 512 && h->sct_W > 4) h->sct_W = 4;
        const char* ve = getenv("PCB_T_VS");
        int vs = h->n > 2048 ? 3 : 2;
        while (vs < h->m - 4 && (size_t)((h->n >> (vs - 1)) - 8) * 32 * h->sizeofT > 16384) vs++;
        if (tvs && h->n <= 2048 && h->n >= 256) vs = tvs < h->m - 4 ? tvs : h->m - 4;
        if (h->n > 2048) { vs = h->m - 6; if (!(wv == 1 || wv == 2 || wv == 4 || wv == 8)) h->sct_W = 4; }
        if (vs < 2) vs = 2;
        if (ve && atoi(ve) >= (h->n > 2048 ? 3 : 2) && atoi(ve) <= h->m - 4) vs = atoi(ve);
        int bps = 0;
        if (sct_occupancy(h->m, h->fp64, exact, h->sct_W, vs, &bps, &h->sct_smem, &h->sct_glob) != cudaSuccess || bps == 0) {
            cudaGetLastError(); h->sct = false;
        } else {
            h->sct_vs = vs; h->sct_bps = bps;
            const char* be = getenv("PCB_T_BPS");                // experiment: fewer resident tiles per SM
            if (be && atoi(be) > 0 && atoi(be) < h->sct_bps) h->sct_bps = atoi(be);
        }
    }
    if (h->sc32) {
        // every resident warp keeps a 64 KB (fp64) / 32 KB (fp32) level buffer that has to stay in L2:
        // measured optimum 8 / 14 warps per SM (more warps spill it to DRAM)
        int cap = sc32_is_big(h->m) ? 9 : h->fp64 ? 8 : 14;   // n > 2048: DRAM-bound, each warp owns n * 32 values of scratch (measured: 4/6/9/12 warps -> 0.54/0.64/0.78/0.72 M frames/s)
        if (const char* cenv = getenv("PCB_S_CAP")) { if (atoi(cenv) > 0) cap = atoi(cenv); }   // experiment: another cap
        if (h->sc32_bps > cap) h->sc32_bps = cap;
        const char* benv = getenv("PCB_S_BPS");         // experiment: fewer resident warps per SM
        if (benv && atoi(benv) > 0 && atoi(benv) < h->sc32_bps) h->sc32_bps = atoi(benv);
    }
    const char* env = getenv("PCB_WARP_SMEM");
    const int budget = env ? atoi(env) : 19200;
    const int mode = h->decoder;
    int ls = 1;
    for (; ls < h->m; ls++) {
        const int tot = warp_layout(h->sizeofT, h->L, h->n, ls, h->stride, mode).total;
        if (tot <= budget) break;
        const int pbytes = h->sizeofT * h->L * ((h->n >> (ls - 1)) - 1);
        if (pbytes <= 4096) break;
    }
    h->ls = ls;
    h->warp_smem = warp_layout(h->sizeofT, h->L, h->n, ls, h->stride, mode).total;
    h->scratch_per_warp = (warp_scratch_bytes(h->sizeofT, h->L, h->n, ls) + 255) & ~(size_t)255;
    int best_w = 0, best_wpb = 1, best_bps = 1;
    const int cands[] = {1, 2, 3, 4, 6, 8};
    for (int wpb : cands) {
        if (wenv && atoi(wenv) != wpb) continue;
        const size_t smem = (size_t)h->warp_smem * wpb;
        if (smem > 227 * 1024) continue;
        int bps = 0;
        if (decode_generic_occupancy(h->L, mode, h->fp64, wpb, smem, &bps) != cudaSuccess) { cudaGetLastError(); continue; }
        if (bps * wpb > best_w) { best_w = bps * wpb; best_wpb = wpb; best_bps = bps; }
    }
    if (best_w == 0) return fail(PCB_EINVAL, "configuration does not fit one SM (warp slice %d B)", h->warp_smem);
    h->wpb = best_wpb;
    h->bps = best_bps;
    return PCB_OK;
}

extern "C" {

const char* pcb_last_error(void) { return g_err; }
const char* pcb_version(void) { return "polarcodes_b200 0.1 (sm_100a)"; }

int pcb_device_count(void) {
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) { cudaGetLastError(); return 0; }
    return c;
}

pcb_handle* pcb_create(int device, int n, int k, int crc, int L, int decoder, int precision) {
    int m = 0;
    while ((1 << m) < n) m++;
    if (n < 2 || n > 65536 || (1 << m) != n) { fail(PCB_EINVAL, "n must be a power of two in [2, 65536]"); return nullptr; }
    if (decoder < 1 || decoder > 3) { fail(PCB_EINVAL, "decoder must be 1 (SC), 2 (SCMinsum) or 3 (SCLMinsum)"); return nullptr; }
    if ((precision & 0xFF) > PCB_PREC_INT8 || precision < 0 || ((precision & 0xFF) == PCB_PREC_INT8 && decoder != 2)) {
        fail(PCB_EINVAL, "precision must be PCB_PREC_STRICT, PCB_PREC_FP32 or (decoder 2 only) PCB_PREC_INT8_Q(q)");
        return nullptr;
    }
    if (decoder != 3) L = 1;   // the CRC still matters to decoders 1, 2: the loop appends it for every decoder (PolarCodes.cpp:136-155)
    if (L < 1 || L > 32 || (L & (L - 1))) { fail(PCB_EINVAL, "L must be a power of two in [1, 32]"); return nullptr; }
    if (crc != 0 && crc != 8 && crc != 16) { fail(PCB_EINVAL, "crc must be 0, 8 or 16"); return nullptr; }
    if (k < 1 || k + crc > n) { fail(PCB_EINVAL, "need 1 <= k and k + crc <= n"); return nullptr; }
    if (crc && k % 8) { fail(PCB_EINVAL, "k must be a multiple of 8 when a crc is used (Resources.cpp:108-138)"); return nullptr; }
    int lb = 0;
    while ((1 << lb) < L) lb++;
    if (decoder == 3 && k + crc < lb) { fail(PCB_EINVAL, "k + crc must be >= log2(L) (the list never fills otherwise)"); return nullptr; }
    if (lb * (m - 1) > 64) { fail(PCB_EINVAL, "log2(L) * (log2(n) - 1) must be <= 64"); return nullptr; }
    int cnt = 0;
    if (cudaGetDeviceCount(&cnt) != cudaSuccess || cnt == 0) {
        cudaGetLastError();
        fail(PCB_ENODEV, "no CUDA device: polarcodes_b200 has no CPU fallback");
        return nullptr;
    }
    if (device < 0 || device >= cnt) { fail(PCB_EINVAL, "device %d out of range (%d devices)", device, cnt); return nullptr; }
    pcb_handle* h = new pcb_handle();
    h->device = device; h->n = n; h->m = m; h->k = k; h->crc = crc; h->L = L; h->decoder = decoder;
    h->precision = precision & 0xFF;
    h->qfrac = (precision >> 8) & 0xF;
    h->fp64 = (decoder == 1) || (decoder == 2 && h->precision == PCB_PREC_STRICT);
    h->sizeofT = h->fp64 ? 8 : 4;
    h->stride = decoder == 3 ? k + crc : n;
    h->EW = (h->stride + 31) / 32;
    auto bail = [&](const char* what, cudaError_t e) -> pcb_handle* {
        fail(PCB_ECUDA, "%s: %s", what, cudaGetErrorString(e));
        pcb_destroy(h);
        return nullptr;
    };
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail("cudaGetDeviceProperties", e);
    h->sm_count = prop.multiProcessorCount;
    if (configure(h) != PCB_OK) { pcb_destroy(h); return nullptr; }
    if ((e = cudaStreamCreateWithFlags(&h->sc, cudaStreamNonBlocking)) != cudaSuccess) return bail("stream", e);
    if ((e = cudaStreamCreateWithFlags(&h->sin, cudaStreamNonBlocking)) != cudaSuccess) return bail("stream", e);
    if ((e = cudaStreamCreateWithFlags(&h->sout, cudaStreamNonBlocking)) != cudaSuccess) return bail("stream", e);
    for (int s = 0; s < 2; s++) {
        if ((e = cudaEventCreateWithFlags(&h->ev_in[s], cudaEventDisableTiming)) != cudaSuccess) return bail("event", e);
        if ((e = cudaEventCreateWithFlags(&h->ev_k[s], cudaEventDisableTiming)) != cudaSuccess) return bail("event", e);
        if ((e = cudaEventCreateWithFlags(&h->ev_out[s], cudaEventDisableTiming)) != cudaSuccess) return bail("event", e);
        if ((e = cudaEventCreate(&h->ev_c0[s])) != cudaSuccess) return bail("event", e);
        if ((e = cudaEventCreate(&h->ev_c1[s])) != cudaSuccess) return bail("event", e);
    }
    if ((e = cudaEventCreateWithFlags(&h->ev_caller, cudaEventDisableTiming)) != cudaSuccess) return bail("event", e);
    if ((e = cudaEventCreate(&h->t0)) != cudaSuccess) return bail("event", e);
    if ((e = cudaEventCreate(&h->t1)) != cudaSuccess) return bail("event", e);
    const int NW = (n + 31) / 32;
    if ((e = cudaMalloc(&h->d_info_mask, NW * 4)) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&h->d_frozen_val, NW * 4)) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&h->d_info_pos, (size_t)n * 2)) != cudaSuccess) return bail("cudaMalloc", e);
    // sized for the full generic grid: pcb_set_code may route a mask with frozen ones to the generic kernel
    const size_t sbytes = h->scratch_per_warp * (size_t)h->sm_count * h->bps * h->wpb;
    if (sbytes && (e = cudaMalloc(&h->d_scratch, sbytes)) != cudaSuccess) return bail("cudaMalloc(scratch)", e);
    if ((e = cudaMalloc(&h->d_counters, 16)) != cudaSuccess) return bail("cudaMalloc", e);
    // set-up for the no-allocation entry point: the static scratch array's address on THIS device (current since cudaSetDevice above)
    if ((e = noalloc_scratch_address(&h->d_noalloc, &h->noalloc_bytes)) != cudaSuccess) return bail("cudaGetSymbolAddress(noalloc scratch)", e);
    return h;
}

void pcb_destroy(pcb_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    cudaFree(h->d_info_mask); cudaFree(h->d_frozen_val); cudaFree(h->d_info_pos); cudaFree(h->d_scratch);
    for (int s = 0; s < 2; s++) {
        cudaFree(h->d_y[s]); cudaFree(h->d_sol[s]); cudaFree(h->d_solb[s]);
        if (h->ev_in[s]) cudaEventDestroy(h->ev_in[s]);
        if (h->ev_c0[s]) cudaEventDestroy(h->ev_c0[s]);
        if (h->ev_c1[s]) cudaEventDestroy(h->ev_c1[s]);
        if (h->ev_k[s]) cudaEventDestroy(h->ev_k[s]);
        if (h->ev_out[s]) cudaEventDestroy(h->ev_out[s]);
    }
    cudaFree(h->d_ymc); cudaFree(h->d_tx); cudaFree(h->d_dec); cudaFree(h->d_counters);
    cudaFree(h->d_eu); cudaFree(h->d_ex); cudaFree(h->d_epu[0]); cudaFree(h->d_epu[1]); cudaFree(h->d_epx[0]); cudaFree(h->d_epx[1]); cudaFree(h->d_redo); cudaFree(h->d_qscratch); cudaFree(h->d_synd); cudaFree(h->d_gcrc); cudaFree(h->d_mc_redo); cudaFree(h->d_xhat); cudaFree(h->d_skip); cudaFree(h->d_y32); cudaFree(h->d_yscr); cudaFree(h->d_yt); cudaFree(h->d_sc32_scratch); cudaFree(h->d_sc32_bits); cudaFree(h->d_sct_ops); cudaFree(h->d_sct_glev); cudaFree(h->d_sc8_ops); cudaFree(h->d_yq); cudaFree(h->d_sc8_glev);
    // ncclCommDestroy may wait for the other ranks of the communicator; a handle is destroyed by its owner alone (the work on
    // its stream has been synchronised above), so the communicator is aborted where the library offers that
    if (h->comm) { NcclApi& nc = nccl_api(); if (nc.CommAbort) nc.CommAbort(h->comm); else nc.CommDestroy(h->comm); }
    cudaFree(h->d_red);
    cudaFreeHost(h->h_stage[0]); cudaFreeHost(h->h_stage[1]); delete h->narrow;
    if (h->ev_caller) cudaEventDestroy(h->ev_caller);
    if (h->t0) cudaEventDestroy(h->t0);
    if (h->t1) cudaEventDestroy(h->t1);
    if (h->sc) cudaStreamDestroy(h->sc);
    if (h->sin) cudaStreamDestroy(h->sin);
    if (h->sout) cudaStreamDestroy(h->sout);
    cudaGetLastError();
    delete h;
}

int pcb_set_code(pcb_handle* h, const int* frozen_mask, double sigma) {
    if (!h || !frozen_mask) return fail(PCB_EINVAL, "pcb_set_code: null argument");
    if (int r = ensure_device(h)) return r;
    const int n = h->n, NW = (n + 31) / 32;
    std::vector<uint32_t> im(NW, 0), fv(NW, 0);
    std::vector<uint16_t> pos;
    for (int i = 0; i < n; i++) {
        if (frozen_mask[i] == -1) { im[i >> 5] |= 1u << (i & 31); pos.push_back((uint16_t)i); }
        else if (frozen_mask[i] == 1) fv[i >> 5] |= 1u << (i & 31);
        else if (frozen_mask[i] != 0) return fail(PCB_EINVAL, "frozen_mask[%d] = %d (want -1, 0 or 1)", i, frozen_mask[i]);
    }
    if (h->decoder == 3 && (int)pos.size() != h->k + h->crc)
        return fail(PCB_EINVAL, "frozen mask has %d information positions, expected k + crc = %d", (int)pos.size(),
                    h->k + h->crc);
    h->ninfo = (int)pos.size();     // decoders 1, 2 decode any mask; pcb_montecarlo / pcb_generate_* want k + crc positions
    pos.resize(n, 0);
    CK(cudaMemcpyAsync(h->d_info_mask, im.data(), NW * 4, cudaMemcpyHostToDevice, h->sc));
    CK(cudaMemcpyAsync(h->d_frozen_val, fv.data(), NW * 4, cudaMemcpyHostToDevice, h->sc));
    CK(cudaMemcpyAsync(h->d_info_pos, pos.data(), (size_t)n * 2, cudaMemcpyHostToDevice, h->sc));
    CK(cudaStreamSynchronize(h->sc));
    h->h_info_w = im; h->h_fv_w = fv;
    if (h->sc8) {
        const int nops = sc8_build_ops(n, h->m, im.data(), fv.data(), nullptr, 0);
        std::vector<uint32_t> ops((size_t)nops + 1);
        sc8_build_ops(n, h->m, im.data(), fv.data(), ops.data(), nops);
        if (h->d_sc8_ops) { CK(cudaFree(h->d_sc8_ops)); h->d_sc8_ops = nullptr; }
        CK(cudaMalloc(&h->d_sc8_ops, ops.size() * 4));
        CK(cudaMemcpyAsync(h->d_sc8_ops, ops.data(), ops.size() * 4, cudaMemcpyHostToDevice, h->sc));
        CK(cudaStreamSynchronize(h->sc));
        h->sc8_nops = nops;
    }
    if (h->sct) {
        // the frozen pattern decides which subtrees need no LLR work at all: compile the pruned tree walk once per code
        int npops = 0;
        const int nops = sct_build_ops(n, h->m, h->sct_W, im.data(), fv.data(), nullptr, 0, nullptr, 0, &npops);
        std::vector<uint32_t> ops((size_t)nops + npops + 1);
        sct_build_ops(n, h->m, h->sct_W, im.data(), fv.data(), ops.data(), nops, ops.data() + nops, npops, nullptr);
        if (h->d_sct_ops) { CK(cudaFree(h->d_sct_ops)); h->d_sct_ops = nullptr; }
        CK(cudaMalloc(&h->d_sct_ops, ops.size() * 4));
        CK(cudaMemcpyAsync(h->d_sct_ops, ops.data(), ops.size() * 4, cudaMemcpyHostToDevice, h->sc));
        CK(cudaStreamSynchronize(h->sc));
        h->sct_npops = npops;
        h->sct_nops = nops;
    }
    if (h->decoder == 3 && h->crc > 0 && h->quad) {
        // The final selection (SCLDecoderMinsum.cpp:442-472) keeps the first slot whose CRC passes:
        // crc(sol[0..k)) == the bits sol[k..k+crc), with sol the information positions of
        // u = x F^(x)m.  Both sides are GF(2)-linear in the codeword estimate x up to the constant
        // the CRC register's initial value contributes, so the test is a syndrome: slot passes
        // iff S x == c0, c0 = f(0), column i of S = f(e_i) ^ f(0) for
        // f(x) = crc(sol(x)[0..k)) ^ (sol(x)[k..k+crc) read MSB first).  u(e_i)_j = 1 iff j is a
        // bitwise subset of i.  The list kernel evaluates S x with 16 parities per word.
        const int K = h->k, C = h->crc, KC = K + C;
        std::vector<int> sol(KC, 0);
        auto fval = [&]() {
            unsigned r = C == 8 ? pcb_crc8(sol.data(), K) : pcb_crc16(sol.data(), K), rx = 0;
            for (int t = 0; t < C; t++) rx = (rx << 1) | (unsigned)sol[K + t];
            return r ^ rx;
        };
        const unsigned c0 = fval();
        std::vector<uint32_t> S((size_t)16 * NW, 0);
        for (int i = 0; i < n; i++) {
            for (int j = 0; j < KC; j++) sol[j] = ((int)pos[j] & i) == (int)pos[j];
            const unsigned col = fval() ^ c0;
            for (int t = 0; t < C; t++)
                if ((col >> t) & 1) S[(size_t)t * NW + (i >> 5)] |= 1u << (i & 31);
        }
        if (!h->d_synd) CK(cudaMalloc(&h->d_synd, (size_t)16 * NW * 4));
        CK(cudaMemcpyAsync(h->d_synd, S.data(), (size_t)16 * NW * 4, cudaMemcpyHostToDevice, h->sc));
        CK(cudaStreamSynchronize(h->sc));   // S is a local: the copy has landed before it goes away; ordered before any decode on h->sc
        h->synd_c0 = c0;
        // The generating prologue of the list kernel (Monte-Carlo mode) computes the CRC the loop appends
        // (PolarCodes.cpp:136-155) the same way: crc(payload) = crc(0) ^ XOR of the columns of the set payload bits.
        const int IW = (KC + 31) / 32;
        std::vector<int> msg(K, 0);
        const unsigned g0 = C == 8 ? pcb_crc8(msg.data(), K) : pcb_crc16(msg.data(), K);
        std::vector<uint32_t> Gm((size_t)16 * IW, 0);
        for (int j = 0; j < K; j++) {
            msg[j] = 1;
            const unsigned col = (C == 8 ? pcb_crc8(msg.data(), K) : pcb_crc16(msg.data(), K)) ^ g0;
            msg[j] = 0;
            for (int t = 0; t < C; t++)
                if ((col >> t) & 1) Gm[(size_t)(j >> 5) * 16 + t] |= 1u << (j & 31);      // [word][crc bit]: 16 consecutive words per payload word
        }
        if (h->d_gcrc) { CK(cudaFree(h->d_gcrc)); h->d_gcrc = nullptr; }
        CK(cudaMalloc(&h->d_gcrc, Gm.size() * 4));
        CK(cudaMemcpyAsync(h->d_gcrc, Gm.data(), Gm.size() * 4, cudaMemcpyHostToDevice, h->sc));
        CK(cudaStreamSynchronize(h->sc));
        h->gcrc_c0 = g0;
    }
    // the tuned list-8 kernel assumes all-zero frozen values (what the reference uses,
    // PolarCodes.cpp:127); a mask with frozen ones runs on the generic kernel
    bool any_fv = false;
    for (uint32_t w : fv) any_fv = any_fv || w != 0;
    h->fast8 = h->f_bps > 0 && !any_fv;
    h->sigma = sigma;
    h->code_set = true;
    return PCB_OK;
}

int pcb_alu_peak(pcb_handle* h, double* lane_ops_per_s) {
    if (!h || !lane_ops_per_s) return fail(PCB_EINVAL, "pcb_alu_peak: null argument");
    int khz = 0;
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, h->device));
    *lane_ops_per_s = (double)h->sm_count * (h->fp64 ? 64.0 : 128.0) * (double)khz * 1e3;
    return PCB_OK;
}

int pcb_info(pcb_handle* h, int* sm_count, int* warps_per_block, int* blocks, int* smem_per_block,
             int* first_smem_layer, uint64_t* kernel_launches) {
    if (!h) return fail(PCB_EINVAL, "pcb_info: null handle");
    if (sm_count) *sm_count = h->sm_count;
    if (h->sc8) {
        if (warps_per_block) *warps_per_block = 1;
        if (blocks) *blocks = h->sm_count * h->sc8_bps;
        if (smem_per_block) *smem_per_block = (int)h->sc8_smem;
        if (first_smem_layer) *first_smem_layer = h->sc8_vs;
        if (kernel_launches) *kernel_launches = h->launches;
        return PCB_OK;
    }
    if (h->sct) {
        if (warps_per_block) *warps_per_block = h->sct_W;
        if (blocks) *blocks = h->sm_count * h->sct_bps;
        if (smem_per_block) *smem_per_block = (int)h->sct_smem;
        if (first_smem_layer) *first_smem_layer = h->sct_vs;
        if (kernel_launches) *kernel_launches = h->launches;
        return PCB_OK;
    }
    const int fw = h->fast8 && h->quad ? scl8q_warps_per_block(h->m, h->L) : 1;
    if (warps_per_block) *warps_per_block = h->fast8 ? fw : h->wpb;
    if (blocks) *blocks = h->fast8 ? fast_blocks(h) / fw : max_blocks(h);
    if (smem_per_block) *smem_per_block = h->fast8 ? h->f_warp_smem * fw : h->warp_smem * h->wpb;
    if (first_smem_layer) *first_smem_layer = h->ls;
    if (kernel_launches) *kernel_launches = h->launches;
    return PCB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ decode plumbing ------
static DecodeArgs make_args(pcb_handle* h, const void* d_y, int y_f32, int* d_sol, uint32_t* d_solb, int B) {
    DecodeArgs a;
    a.n = h->n; a.m = h->m; a.k = h->k; a.crc = h->decoder == 3 ? h->crc : 0; a.B = B;   // the SC decoders never look at the CRC
    a.first_smem_layer = h->ls; a.warp_smem_bytes = h->warp_smem; a.stride = h->stride;
    a.y = d_y; a.y_f32 = y_f32; a.y64_tail = nullptr; a.y64_first = 0;
    a.info_mask = h->d_info_mask; a.frozen_val = h->d_frozen_val; a.info_pos = h->d_info_pos;
    a.sol = d_sol; a.sol_bits = d_solb;
    a.scratch = h->d_scratch; a.scratch_per_warp = h->scratch_per_warp;
    a.sigma = h->sigma;
    a.redo_count = nullptr; a.redo_list = nullptr; a.use_redo_list = 0;
    return a;
}

static cudaError_t ensure_redo(pcb_handle* h, size_t B) {
    if (h->redo_cap >= B + 2) return cudaSuccess;
    cudaFree(h->d_redo); h->d_redo = nullptr; h->redo_cap = 0;     // (re)allocation synchronises the device: safe, and rare
    cudaError_t e0 = cudaMalloc(&h->d_redo, (B + 2) * sizeof(int));
    if (e0 != cudaSuccess) return e0;
    h->redo_cap = B + 2;
    return cudaSuccess;
}

static cudaError_t ensure_xhat(pcb_handle* h, size_t B) {
    if (h->xhat_cap >= B) return cudaSuccess;
    cudaFree(h->d_xhat); h->d_xhat = nullptr; h->xhat_cap = 0;
    cudaError_t e0 = cudaMalloc(&h->d_xhat, B * (size_t)h->L * (h->n / 32) * sizeof(uint32_t));
    if (e0 != cudaSuccess) return e0;
    h->xhat_cap = B;
    return cudaSuccess;
}

static SelectArgs select_args(pcb_handle* h, const DecodeArgs& a) {
    SelectArgs s;
    s.xhat = h->d_xhat; s.B = a.B; s.L = h->L; s.n = h->n; s.k = h->k; s.crc = h->crc; s.stride = h->stride;
    s.info_mask = h->d_info_mask; s.synd = h->d_synd; s.synd_c0 = h->synd_c0;
    s.sol = a.sol; s.sol_bits = a.sol_bits;
    s.count = 0; s.skip = nullptr; s.counters = nullptr; s.seed = 0; s.first_frame = 0;
    return s;
}

// prechecked: the producer of the (float) batch has already zeroed the redo header and listed the frames
// outside the tuned list kernel's safe input range (the on-device generator does): no pre-pass
static cudaError_t launch(pcb_handle* h, const DecodeArgs& a, cudaStream_t st, bool prechecked = false) {
    h->launches++;
    if (h->fast8) {
        cudaError_t e = ensure_redo(h, (size_t)a.B);
        if (e != cudaSuccess) return e;
        // tuned kernel first; frames outside its safe input range come back in the redo list and
        // are decoded by the generic kernel right behind it (normally an empty launch)
        DecodeArgs f = a;
        f.redo_count = h->d_redo; f.redo_list = h->d_redo + 2; f.use_redo_list = 0;
        if (!prechecked) {
            e = cudaMemsetAsync(h->d_redo, 0, 2 * sizeof(int), st);
            if (e != cudaSuccess) return e;
        }
        if (h->quad) {
            // The decoder works in float (SCLDecoderMinsum.cpp:406-408 narrows every value first).  The warp that
            // decodes a quad narrows its doubles itself (into the block's rows of d_yscr, which stay in L2) and
            // range-checks its frames for the redo list: no pre-pass over the batch, no float copy in HBM.
            // PCB_SEPARATE_PREPASS=1 keeps the old HBM-bound pre-pass kernel (A/B).
            const float* y32 = reinterpret_cast<const float*>(a.y);
            const double* y64 = nullptr;
            int nf = 0;                      // frames below nf are float already
            if (!f.y_f32) { y64 = reinterpret_cast<const double*>(a.y); y32 = nullptr; }
            else if (h->mix_tail) { y64 = h->mix_tail; nf = h->mix_nf; }   // host pipeline: the tail of the chunk came over as doubles
            h->mix_tail = nullptr;
            const int fpw = scl8q_frames_per_warp(h->L);
            int check = prechecked ? 0 : 1;
            static const bool separate = getenv("PCB_SEPARATE_PREPASS") != nullptr;
            if (separate && !prechecked) {
                float* dst = const_cast<float*>(y32);
                if (!f.y_f32) {
                    const size_t cnt = (size_t)a.B * h->n;
                    if (h->y32_cap < cnt) {
                        cudaFree(h->d_y32); h->d_y32 = nullptr; h->y32_cap = 0;
                        cudaError_t e1 = cudaMalloc(&h->d_y32, cnt * sizeof(float));
                        if (e1 != cudaSuccess) return e1;
                        h->y32_cap = cnt;
                    }
                    dst = h->d_y32;
                }
                e = launch_scl8_prepass(y64, dst, h->n, a.B, nf, f.redo_count, f.redo_list, h->sm_count, st);
                if (e != cudaSuccess) return e;
                h->launches++;
                y32 = dst; y64 = nullptr; check = 0;
            }
            if (y64 && !h->d_yscr) {
                cudaError_t e1 = cudaMalloc(&h->d_yscr, (size_t)fast_blocks(h) * fpw * h->n * sizeof(float));
                if (e1 != cudaSuccess) return e1;
            }
            f.y = y32; f.y_f32 = 1;
            f.y64_tail = y64; f.y64_first = nf;      // the generic kernel finds the redo frames where the list kernel found them
            const int nq = (a.B + fpw - 1) / fpw;
            const size_t spb = scl8q_scratch_per_block(h->m, h->L);
            if (spb && !h->d_qscratch) {
                cudaError_t e1 = cudaMalloc(&h->d_qscratch, spb * (size_t)fast_blocks(h));
                if (e1 != cudaSuccess) return e1;
            }
            e = ensure_xhat(h, (size_t)a.B);
            if (e != cudaSuccess) return e;
            e = launch_decode_scl8q(f, h->L, h->h_info_w.data(), h->d_xhat, h->d_qscratch, h->d_redo + 1, y64, nf, h->d_yscr, check,
                                    nullptr, nq < fast_blocks(h) ? nq : fast_blocks(h), st);
            if (e != cudaSuccess) return e;
            e = launch_scl_select(select_args(h, f), h->sm_count, st);     // the generic kernel behind it overrides the redo frames
            h->launches++;
        } else {
            return cudaErrorInvalidValue;
        }
        if (e != cudaSuccess) return e;
        f.use_redo_list = 1;
        h->launches++;
        return launch_decode_generic(f, h->L, h->decoder, h->fp64, max_blocks(h), h->wpb, st);
    }
    if (h->sc8) {
        const size_t need = sc8_yq_bytes(h->n, a.B);
        if (h->yq_cap < need) {
            cudaFree(h->d_yq); h->d_yq = nullptr; h->yq_cap = 0;
            cudaError_t e0 = cudaMalloc(&h->d_yq, need);
            if (e0 != cudaSuccess) return e0;
            h->yq_cap = need;
        }
        if (!a.sol_bits && h->sc32_bits_cap < (size_t)a.B * (h->n / 32)) {
            cudaFree(h->d_sc32_bits); h->d_sc32_bits = nullptr; h->sc32_bits_cap = 0;
            cudaError_t e0 = cudaMalloc(&h->d_sc32_bits, (size_t)a.B * (h->n / 32) * 4);
            if (e0 != cudaSuccess) return e0;
            h->sc32_bits_cap = (size_t)a.B * (h->n / 32);
        }
        uint32_t* bits = a.sol_bits ? a.sol_bits : h->d_sc32_bits;
        h->launches++;                        // quantiser + decoder
        if (!h->d_sc8_glev) {
            cudaError_t e0 = cudaMalloc(&h->d_sc8_glev, sc8_glev_bytes(h->n, h->sm_count * h->sc8_bps, h->sc8_vs));
            if (e0 != cudaSuccess) return e0;
        }
        cudaError_t e = launch_decode_sc8(a, h->qfrac, h->d_sc8_ops, h->sc8_nops, h->d_yq, h->d_sc8_glev, bits, h->sm_count * h->sc8_bps, h->sm_count, h->sc8_vs, st);
        if (e != cudaSuccess) return e;
        if (a.sol) return launch_sc_expand(bits, a.sol, (size_t)a.B * h->n, h->sm_count, st);
        return cudaSuccess;
    }
    if (h->sct) {
        cudaError_t e = ensure_redo(h, (size_t)a.B);
        if (e != cudaSuccess) return e;
        if (!a.sol_bits && h->sc32_bits_cap < (size_t)a.B * (h->n / 32)) {
            cudaFree(h->d_sc32_bits); h->d_sc32_bits = nullptr; h->sc32_bits_cap = 0;
            cudaError_t e0 = cudaMalloc(&h->d_sc32_bits, (size_t)a.B * (h->n / 32) * 4);
            if (e0 != cudaSuccess) return e0;
            h->sc32_bits_cap = (size_t)a.B * (h->n / 32);
        }
        e = cudaMemsetAsync(h->d_redo, 0, 2 * sizeof(int), st);
        if (e != cudaSuccess) return e;
        // frames in which an all-information subtree saw a zero (or vanishing) LLR come back on the redo list and are
        // decoded by the generic kernel right behind (normally an empty launch)
        DecodeArgs f = a;
        f.redo_count = h->d_redo; f.redo_list = h->d_redo + 2; f.use_redo_list = 0;
        uint32_t* bits = a.sol_bits ? a.sol_bits : h->d_sc32_bits;
        if (h->sct_glob && !h->d_sct_glev) {
            cudaError_t e0 = cudaMalloc(&h->d_sct_glev, h->sct_glob * (size_t)h->sm_count * h->sct_bps);
            if (e0 != cudaSuccess) return e0;
        }
        e = launch_decode_sct(f, h->d_sct_ops, h->sct_nops, h->d_sct_ops + h->sct_nops, h->sct_npops, bits, h->d_sct_glev, h->fp64, h->decoder == 1,
                              h->sct_W, h->sct_vs, h->sm_count * h->sct_bps, st);
        if (e != cudaSuccess) return e;
        if (a.sol) {
            e = launch_sc_expand(bits, a.sol, (size_t)a.B * h->n, h->sm_count, st);
            if (e != cudaSuccess) return e;
        }
        f.use_redo_list = 1;
        h->launches++;
        return launch_decode_generic(f, h->L, h->decoder, h->fp64, max_blocks(h), h->wpb, st);
    }
    if (h->sc32) {
        const size_t need = sc32_yt_bytes(h->n, a.B, h->fp64);
        if (h->yt_cap < need) {               // (re)allocation synchronises the device: safe, and rare
            cudaFree(h->d_yt); h->d_yt = nullptr; h->yt_cap = 0;
            cudaError_t e0 = cudaMalloc(&h->d_yt, need);
            if (e0 != cudaSuccess) return e0;
            h->yt_cap = need;
        }
        if (!h->d_sc32_scratch) {
            const int nb = h->sm_count * h->sc32_bps;
            cudaError_t e0 = cudaMalloc(&h->d_sc32_scratch, sc32_is_big(h->m) ? sc32_big_scratch_bytes(h->n, nb, h->fp64)
                                                                              : sc32_scratch_bytes(nb, h->fp64));
            if (e0 != cudaSuccess) return e0;
        }
        if (!a.sol_bits && h->sc32_bits_cap < (size_t)a.B * (h->n / 32)) {
            cudaFree(h->d_sc32_bits); h->d_sc32_bits = nullptr; h->sc32_bits_cap = 0;
            cudaError_t e0 = cudaMalloc(&h->d_sc32_bits, (size_t)a.B * (h->n / 32) * 4);
            if (e0 != cudaSuccess) return e0;
            h->sc32_bits_cap = (size_t)a.B * (h->n / 32);
        }
        h->launches++;                        // transposition + decoder
        return launch_decode_sc32(a, h->h_info_w.data(), h->h_fv_w.data(), h->d_yt, h->d_sc32_scratch, h->d_sc32_bits, h->fp64, h->decoder == 1,
                                  h->sm_count * h->sc32_bps, h->sm_count, st);
    }
    int blocks = (a.B + h->wpb - 1) / h->wpb;
    if (blocks > max_blocks(h)) blocks = max_blocks(h);
    return launch_decode_generic(a, h->L, h->decoder, h->fp64, blocks, h->wpb, st);
}

static int ensure_slots(pcb_handle* h, size_t frames, size_t ybytes_per_frame) {
    if (h->slot_frames >= frames) return PCB_OK;
    for (int s = 0; s < 2; s++) {
        cudaFree(h->d_y[s]); cudaFree(h->d_sol[s]); cudaFree(h->d_solb[s]);
        h->d_y[s] = nullptr; h->d_sol[s] = nullptr; h->d_solb[s] = nullptr;
    }
    h->slot_frames = 0;
    for (int s = 0; s < 2; s++) {
        // 12 bytes per value: a chunk of the host pipeline may arrive as floats (narrowed on the host)
        // followed by a tail of doubles, [frames*n floats][frames*n doubles]
        CK(cudaMalloc(&h->d_y[s], frames * (size_t)h->n * 12));
        CK(cudaMalloc(&h->d_sol[s], frames * (size_t)h->stride * 4));
        CK(cudaMalloc(&h->d_solb[s], frames * (size_t)h->EW * 4));
    }
    (void)ybytes_per_frame;
    h->slot_frames = frames;
    return PCB_OK;
}

// host buffers -> chunked, double-buffered pipeline: H2D on `sin`, kernel on `sc`, D2H on `sout`
static int decode_host(pcb_handle* h, const void* y, int y_f32, int* sol, uint32_t* solb, int B) {
    if (!h || !y || (!sol && !solb)) return fail(PCB_EINVAL, "pcb_decode_batch: null argument");
    if (!h->code_set) return fail(PCB_ESTATE, "pcb_decode_batch: call pcb_set_code first");
    if (B < 0) return fail(PCB_EINVAL, "pcb_decode_batch: B < 0");
    if (B == 0) return PCB_OK;
    if (int r = ensure_device(h)) return r;
    const size_t ybpf = (size_t)h->n * (y_f32 ? 4 : 8);
    size_t chunk = ((size_t)48 << 20) / ((size_t)h->n * 8);
    const char* cenv = getenv("PCB_CHUNK_FRAMES");
    if (cenv) chunk = (size_t)atol(cenv);
    if (chunk < 1) chunk = 1;
    if (chunk > (size_t)B) chunk = (size_t)B;
    if (int r = ensure_slots(h, chunk, ybpf)) return r;
    // float decoders fed with doubles: narrow on the host cores, copy 4 bytes per value
    const bool narrow = !y_f32 && !h->fp64 && !h->sc8 && !getenv("PCB_NO_HOST_NARROW");   // (the 8-bit mode quantises the doubles themselves)
    const bool mixed = narrow && h->fast8 && h->quad && !getenv("PCB_NO_MIXED_INPUT");
    const double* mix_tail = nullptr;
    int mix_nf = 0;
    auto t_iter = std::chrono::steady_clock::now();
    if (narrow) {
        if (h->stage_frames < chunk) {
            cudaFreeHost(h->h_stage[0]); cudaFreeHost(h->h_stage[1]);
            h->h_stage[0] = h->h_stage[1] = nullptr; h->stage_frames = 0;
            for (int s = 0; s < 2; s++) CK(cudaHostAlloc(&h->h_stage[s], chunk * (size_t)h->n * 4, cudaHostAllocDefault));
            h->stage_frames = chunk;
        }
        if (!h->narrow) {
            h->narrow = new HostNarrow();
            const char* lw = getenv("LOCAL_WORLD_SIZE");
            h->shared_host = lw && atoi(lw) > 1;
            if (h->shared_host) h->narrow_frac = 0.3;
        }
        h->narrow->start();
    }
    int c = 0;
    for (size_t f0 = 0; f0 < (size_t)B; f0 += chunk, c++) {
        const int s = c & 1;
        const int nb = (int)((size_t)B - f0 < chunk ? (size_t)B - f0 : chunk);
        if (c >= 2) CK(cudaStreamWaitEvent(h->sin, h->ev_k[s], 0));      // slot input consumed
        if (narrow) {
            if (c >= 2) {
                CK(cudaEventSynchronize(h->ev_in[s]));                   // staging slot copied out
                float ms = 0.f;                                          // H2D rate of the chunk that used this slot
                if (h->copy_bytes[s] && cudaEventElapsedTime(&ms, h->ev_c0[s], h->ev_c1[s]) == cudaSuccess && ms > 0.f) {
                    const double r = (double)h->copy_bytes[s] / (ms * 1e-3);
                    h->rate_h2d = h->rate_h2d > 0 ? 0.5 * h->rate_h2d + 0.5 * r : r;
                }
            }
            // The list-8 pre-pass can convert on the device, so the chunk is split: the head is
            // narrowed here (4 bytes per value over PCIe), the tail goes over as doubles while the
            // host cores work.  A process that has the host to itself uses the split at which cores
            // and link finish together, 8R / (W + 4R) for R values/s narrowed and W bytes/s copied
            // (both measured per chunk).  Ranks that share a host also share its memory system, R
            // and W then depend on what the others do: there the share of the head follows the
            // measured chunk rate (every three chunks it moves by 0.08, and turns round when the
            // rate dropped), starting low.
            int nf = nb;
            if (mixed) {
                const auto now = std::chrono::steady_clock::now();
                if (!h->shared_host) {
                    if (h->rate_cpu > 0 && h->rate_h2d > 0) h->narrow_frac = 8.0 * h->rate_cpu / (h->rate_h2d + 4.0 * h->rate_cpu);
                    if (h->narrow_frac < 0.1) h->narrow_frac = 0.1;      // keeps the host rate measured
                    if (h->narrow_frac > 1.0) h->narrow_frac = 1.0;
                } else if (c >= 2) {                                     // steady state: this loop paces the pipeline
                    h->pc_time += std::chrono::duration<double>(now - t_iter).count();
                    h->pc_frames += (double)chunk;
                    if (++h->pc_chunks == 3) {
                        const double tp = h->pc_frames / h->pc_time;
                        if (h->pc_prev > 0 && tp < 0.99 * h->pc_prev) h->pc_dir = -h->pc_dir;
                        h->pc_prev = tp;
                        h->narrow_frac += 0.08 * h->pc_dir;
                        if (h->narrow_frac >= 1.0) { h->narrow_frac = 1.0; h->pc_dir = -1.0; }
                        if (h->narrow_frac <= 0.0) { h->narrow_frac = 0.0; h->pc_dir = 1.0; }
                        h->pc_frames = h->pc_time = 0; h->pc_chunks = 0;
                    }
                }
                t_iter = now;
                const char* fe = getenv("PCB_NARROW_FRAC");              // experiment: fixed split
                const double frac = (fe && *fe) ? atof(fe) : h->narrow_frac;
                nf = ((int)(frac * nb) + 3) & ~3;
                if (nf > nb) nf = nb;
            }
            const int nd = nb - nf;
            double* d_tail = reinterpret_cast<double*>(reinterpret_cast<char*>(h->d_y[s]) + chunk * (size_t)h->n * 4);
            CK(cudaEventRecord(h->ev_c0[s], h->sin));
            if (nd) CK(cudaMemcpyAsync(d_tail, (const double*)y + (f0 + nf) * h->n, (size_t)nd * h->n * 8, cudaMemcpyHostToDevice, h->sin));
            CK(cudaEventRecord(h->ev_c1[s], h->sin));                    // the tail copy alone times the link (no host work inside)
            h->copy_bytes[s] = (size_t)nd * h->n * 8;
            const auto t0 = std::chrono::steady_clock::now();
            if (nf) h->narrow->run((const double*)y + f0 * h->n, h->h_stage[s], (size_t)nf * h->n);
            const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (nf && dt > 0) { const double r = (double)nf * h->n / dt; h->rate_cpu = h->rate_cpu > 0 ? 0.5 * h->rate_cpu + 0.5 * r : r; }
            if (nf) CK(cudaMemcpyAsync(h->d_y[s], h->h_stage[s], (size_t)nf * h->n * 4, cudaMemcpyHostToDevice, h->sin));
            mix_tail = nd ? d_tail : nullptr; mix_nf = nf;
        } else {
            CK(cudaMemcpyAsync(h->d_y[s], (const char*)y + f0 * ybpf, (size_t)nb * ybpf, cudaMemcpyHostToDevice, h->sin));
        }
        CK(cudaEventRecord(h->ev_in[s], h->sin));
        CK(cudaStreamWaitEvent(h->sc, h->ev_in[s], 0));
        if (c >= 2) CK(cudaStreamWaitEvent(h->sc, h->ev_out[s], 0));     // slot output drained
        DecodeArgs a = make_args(h, h->d_y[s], y_f32 || narrow, sol ? h->d_sol[s] : nullptr, solb ? h->d_solb[s] : nullptr, nb);
        h->mix_tail = mix_tail; h->mix_nf = mix_nf; mix_tail = nullptr;
        CK(launch(h, a, h->sc));
        CK(cudaEventRecord(h->ev_k[s], h->sc));
        CK(cudaStreamWaitEvent(h->sout, h->ev_k[s], 0));
        if (sol)
            CK(cudaMemcpyAsync(sol + f0 * h->stride, h->d_sol[s], (size_t)nb * h->stride * 4, cudaMemcpyDeviceToHost, h->sout));
        if (solb)
            CK(cudaMemcpyAsync(solb + f0 * h->EW, h->d_solb[s], (size_t)nb * h->EW * 4, cudaMemcpyDeviceToHost, h->sout));
        CK(cudaEventRecord(h->ev_out[s], h->sout));
    }
    CK(cudaStreamSynchronize(h->sout));
    CK(cudaStreamSynchronize(h->sc));
    CK(cudaStreamSynchronize(h->sin));
    if (narrow && getenv("PCB_DEBUG_PIPE"))
        fprintf(stderr, "polar_b200 host pipeline: %d threads, %.2f Gvalues/s narrowed, %.1f GB/s H2D, head share %.2f\n",
                h->narrow->T, h->rate_cpu / 1e9, h->rate_h2d / 1e9, h->narrow_frac);
    return PCB_OK;
}

// The device-pointer entry points run on the handle's own non-blocking stream, which does not wait for the legacy default
// stream the caller's buffers are usually produced on (torch's current stream unless the caller changed it): order the
// call after everything already queued there.  (A caller with its own non-blocking streams synchronises them itself.)
static cudaError_t after_callers_stream(pcb_handle* h) {
    const cudaError_t e = cudaEventRecord(h->ev_caller, cudaStreamLegacy);
    return e != cudaSuccess ? e : cudaStreamWaitEvent(h->sc, h->ev_caller, 0);
}

extern "C" {

int pcb_decode_batch(pcb_handle* h, const double* y, int* sol, int B) { return decode_host(h, y, 0, sol, nullptr, B); }
int pcb_decode_batch_f32(pcb_handle* h, const float* y, int* sol, int B) { return decode_host(h, y, 1, sol, nullptr, B); }
int pcb_decode_batch_f32_packed(pcb_handle* h, const float* y, uint32_t* sol_bits, int B) {
    return decode_host(h, y, 1, nullptr, sol_bits, B);
}

int pcb_decode_batch_dev(pcb_handle* h, const void* y_dev, int y_is_f32, uint32_t* sol_bits_dev, int B, float* ms) {
    if (!h || !y_dev || !sol_bits_dev) return fail(PCB_EINVAL, "pcb_decode_batch_dev: null argument");
    if (!h->code_set) return fail(PCB_ESTATE, "pcb_decode_batch_dev: call pcb_set_code first");
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "B < 0");
    if (int r = ensure_device(h)) return r;
    DecodeArgs a = make_args(h, y_dev, y_is_f32, nullptr, sol_bits_dev, B);
    CK(after_callers_stream(h));
    CK(cudaEventRecord(h->t0, h->sc));
    CK(launch(h, a, h->sc));
    CK(cudaEventRecord(h->t1, h->sc));
    CK(cudaStreamSynchronize(h->sc));
    if (ms) CK(cudaEventElapsedTime(ms, h->t0, h->t1));
    return PCB_OK;
}

// No-allocation variant of pcb_decode_batch_dev: the generic (reference-literal) kernel with its per-warp LLR scratch
// in the static device array of decode_generic.cu instead of the handle's cudaMalloc'ed one.  Nothing here allocates:
// no cudaMalloc / cudaFree, no lazily grown handle buffer (the tuned kernels' redo list, codeword estimates and
// transposed batch are what pcb_decode_batch_dev grows on demand, so this call does not use them).
int pcb_decode_batch_dev_noalloc(pcb_handle* h, const void* y_dev, int y_is_f32, uint32_t* sol_bits_dev, int B) {
    if (!h || !y_dev || !sol_bits_dev) return fail(PCB_EINVAL, "pcb_decode_batch_dev_noalloc: null argument");
    if (!h->code_set) return fail(PCB_ESTATE, "pcb_decode_batch_dev_noalloc: call pcb_set_code first");
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "pcb_decode_batch_dev_noalloc: B < 0");
    if (!h->d_noalloc) return fail(PCB_ESTATE, "pcb_decode_batch_dev_noalloc: the static scratch array has no address on device %d", h->device);
    if (int r = ensure_device(h)) return r;
    // one warp decodes one frame at a time and owns scratch_per_warp bytes of the array: as many blocks as fit
    const size_t per_block = h->scratch_per_warp * (size_t)h->wpb;
    if (per_block > h->noalloc_bytes)
        return fail(PCB_EINVAL,
                    "pcb_decode_batch_dev_noalloc: n = %d, L = %d needs %zu bytes of LLR scratch per block of %d warps, the static "
                    "array holds %zu (PCB_NOALLOC_SCRATCH_BYTES); use pcb_decode_batch_dev",
                    h->n, h->L, per_block, h->wpb, h->noalloc_bytes);
    if (B > PCB_NOALLOC_MAX_FRAMES)
        return fail(PCB_EINVAL, "pcb_decode_batch_dev_noalloc: B = %d exceeds the limit of %d frames per call", B, PCB_NOALLOC_MAX_FRAMES);
    int blocks = (B + h->wpb - 1) / h->wpb;
    const int resident = h->sm_count * h->bps;
    if (blocks > resident) blocks = resident;
    if (per_block && (size_t)blocks > h->noalloc_bytes / per_block) blocks = (int)(h->noalloc_bytes / per_block);
    DecodeArgs a = make_args(h, y_dev, y_is_f32, nullptr, sol_bits_dev, B);
    a.scratch = h->d_noalloc;
    CK(after_callers_stream(h));
    CK(launch_decode_generic(a, h->L, h->decoder, h->fp64, blocks, h->wpb, h->sc));
    h->launches++;
    CK(cudaStreamSynchronize(h->sc));
    return PCB_OK;
}

int pcb_encode_batch(pcb_handle* h, const int* u, int* x, int with_shuffle, int B) {
    if (!h || !u || !x) return fail(PCB_EINVAL, "pcb_encode_batch: null argument");
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "B < 0");
    if (int r = ensure_device(h)) return r;
    size_t chunk = ((size_t)64 << 20) / ((size_t)h->n * 4);
    if (chunk < 1) chunk = 1;
    if (chunk > (size_t)B) chunk = (size_t)B;
    if (h->enc_frames < chunk) {
        cudaFree(h->d_eu); cudaFree(h->d_ex);
        h->d_eu = h->d_ex = nullptr; h->enc_frames = 0;
        CK(cudaMalloc(&h->d_eu, chunk * (size_t)h->n * 4));
        CK(cudaMalloc(&h->d_ex, chunk * (size_t)h->n * 4));
        h->enc_frames = chunk;
    }
    for (size_t f0 = 0; f0 < (size_t)B; f0 += chunk) {
        const int nb = (int)((size_t)B - f0 < chunk ? (size_t)B - f0 : chunk);
        CK(cudaMemcpyAsync(h->d_eu, u + f0 * h->n, (size_t)nb * h->n * 4, cudaMemcpyHostToDevice, h->sc));
        CK(launch_encode(h->d_eu, h->d_ex, h->n, h->m, with_shuffle != 0, nb, h->sm_count, h->sc));
        h->launches++;
        CK(cudaMemcpyAsync(x + f0 * h->n, h->d_ex, (size_t)nb * h->n * 4, cudaMemcpyDeviceToHost, h->sc));
        CK(cudaStreamSynchronize(h->sc));
    }
    return PCB_OK;
}

// Packed variants of pcb_encode_batch (Encoder.cpp:77-83): 32 bits per word in and out.
int pcb_encode_batch_dev(pcb_handle* h, const uint32_t* u_bits_dev, uint32_t* x_bits_dev, int with_shuffle, int B, float* ms) {
    if (!h || !u_bits_dev || !x_bits_dev) return fail(PCB_EINVAL, "pcb_encode_batch_dev: null argument");
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "B < 0");
    if (int r = ensure_device(h)) return r;
    CK(after_callers_stream(h));
    CK(cudaEventRecord(h->t0, h->sc));
    CK(launch_encode_packed(u_bits_dev, x_bits_dev, h->n, h->m, with_shuffle != 0, B, h->sm_count, h->sc));
    CK(cudaEventRecord(h->t1, h->sc));
    h->launches++;
    CK(cudaStreamSynchronize(h->sc));
    if (ms) CK(cudaEventElapsedTime(ms, h->t0, h->t1));
    return PCB_OK;
}

int pcb_encode_batch_packed(pcb_handle* h, const uint32_t* u_bits, uint32_t* x_bits, int with_shuffle, int B) {
    if (!h || !u_bits || !x_bits) return fail(PCB_EINVAL, "pcb_encode_batch_packed: null argument");
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "B < 0");
    if (int r = ensure_device(h)) return r;
    const size_t NW = h->n >= 32 ? h->n / 32 : 1;
    size_t chunk = ((size_t)16 << 20) / (NW * 4);                      // two slots of 16 MiB each way
    if (chunk > (size_t)B) chunk = (size_t)B;
    if (h->encp_frames < chunk) {
        for (int s2 = 0; s2 < 2; s2++) { cudaFree(h->d_epu[s2]); cudaFree(h->d_epx[s2]); h->d_epu[s2] = h->d_epx[s2] = nullptr; }
        h->encp_frames = 0;
        for (int s2 = 0; s2 < 2; s2++) { CK(cudaMalloc(&h->d_epu[s2], chunk * NW * 4)); CK(cudaMalloc(&h->d_epx[s2], chunk * NW * 4)); }
        h->encp_frames = chunk;
    }
    // copy in / encode / copy out of a chunk are ordered on one stream; a slot is reused two chunks later, behind its copy out
    int c = 0;
    for (size_t f0 = 0; f0 < (size_t)B; f0 += chunk, c++) {
        const int s2 = c & 1;
        const int nb = (int)((size_t)B - f0 < chunk ? (size_t)B - f0 : chunk);
        CK(cudaMemcpyAsync(h->d_epu[s2], u_bits + f0 * NW, (size_t)nb * NW * 4, cudaMemcpyHostToDevice, h->sc));
        CK(launch_encode_packed(h->d_epu[s2], h->d_epx[s2], h->n, h->m, with_shuffle != 0, nb, h->sm_count, h->sc));
        h->launches++;
        CK(cudaMemcpyAsync(x_bits + f0 * NW, h->d_epx[s2], (size_t)nb * NW * 4, cudaMemcpyDeviceToHost, h->sc));
    }
    CK(cudaStreamSynchronize(h->sc));
    return PCB_OK;
}

static int ensure_mc(pcb_handle* h, size_t frames) {
    if (h->mc_frames >= frames) return PCB_OK;
    cudaFree(h->d_ymc); cudaFree(h->d_tx); cudaFree(h->d_dec);
    h->d_ymc = nullptr; h->d_tx = h->d_dec = nullptr; h->mc_frames = 0;
    CK(cudaMalloc(&h->d_ymc, frames * (size_t)h->n * 4));
    CK(cudaMalloc(&h->d_tx, frames * (size_t)h->EW * 4));
    CK(cudaMalloc(&h->d_dec, frames * (size_t)h->EW * 4));
    h->mc_frames = frames;
    return PCB_OK;
}

// frame generation scatters k + crc payload bits (PolarCodes.cpp:133-156): the mask must have as many positions
static int check_gen(const pcb_handle* h, const char* who) {
    if (h->ninfo != h->k + h->crc)
        return fail(PCB_ESTATE, "%s: the frozen mask has %d information positions, frame generation needs k + crc = %d", who,
                    h->ninfo, h->k + h->crc);
    return PCB_OK;
}

static GenArgs gen_args(pcb_handle* h, double snr_db, uint64_t seed, uint64_t first, int B) {
    GenArgs g;
    g.n = h->n; g.m = h->m; g.k = h->k; g.crc = h->crc; g.B = B;
    g.with_shuffle = h->decoder != 3;
    g.tx_is_u = h->decoder != 3;
    g.tx_words = h->EW;
    g.sigma = (float)pcb_sigma(h->n, h->k, snr_db);
    g.seed = seed; g.first_frame = first;
    g.info_mask = h->d_info_mask; g.frozen_val = h->d_frozen_val; g.info_pos = h->d_info_pos;
    g.y = h->d_ymc; g.tx_bits = h->d_tx;
    g.redo_count = nullptr; g.redo_list = nullptr;
    return g;
}

static size_t mc_batch(pcb_handle* h) {
    size_t b = ((size_t)1 << 30) / ((size_t)h->n * 4);     // 1 GiB of channel values: many waves of the resident grid per launch
    const char* env = getenv("PCB_MC_BATCH");
    if (env) b = (size_t)atol(env);
    const size_t minb = h->fast8 ? (size_t)fast_blocks(h) * (h->quad ? scl8q_frames_per_warp(h->L) : 1) : (size_t)max_blocks(h) * h->wpb;
    if (b < minb) b = minb;
    if ((h->sc32 || h->sct) && !env) {
        // lane-per-frame kernels walk the tiles of 32 frames with a static stride: whole waves of
        // the resident grid (one tile per warp and wave), at least one (n = 32768: 42 624 frames)
        const size_t wave = (size_t)h->sm_count * (h->sct ? h->sct_bps : h->sc32_bps) * 32;
        b = (b + wave - 1) / wave * wave;
    }
    return b;
}

int pcb_montecarlo(pcb_handle* h, double snr_db, uint64_t frames, uint64_t seed, uint64_t first_frame_id,
                   uint64_t* frames_done, uint64_t* frame_errors, uint64_t* bit_errors, double* seconds) {
    if (!h) return fail(PCB_EINVAL, "pcb_montecarlo: null handle");
    if (!h->code_set) return fail(PCB_ESTATE, "pcb_montecarlo: call pcb_set_code first");
    if (int r = check_gen(h, "pcb_montecarlo")) return r;
    if (int r = ensure_device(h)) return r;
    size_t batch = mc_batch(h);
    if (h->fast8 && h->quad && getenv("PCB_MC_FUSED") && !getenv("PCB_MC_UNFUSED") && !getenv("PCB_MC_BATCH")) batch = (size_t)1 << 21;   // fused: no channel values in HBM; the launch is the unit (its codeword estimates are L n / 8 bytes per frame)
    if (batch > frames) batch = (size_t)frames;
    if (frames == 0) return PCB_OK;
    CK(cudaMemsetAsync(h->d_counters, 0, 16, h->sc));
    double secs = 0;
    const int nbits = h->decoder == 3 ? h->k : h->n;
    const double mc_sigma = pcb_sigma(h->n, h->k, snr_db);   // exact SC scales by 2/sigma^2 (SCDecoder.cpp:15-17); the handle's sigma stays the caller's
    // generate_kernel -> decoder -> count_errors_kernel, three launches over a float batch in HBM.  For the list decoder
    // on the tuned kernel there is also a one-kernel form: the warp that decodes a quad generates its frames into its
    // rows of d_yscr (they never reach HBM) and scl_select_kernel counts.
    // Measured (config C, crc 16): the generating prologue costs 13.7 % more instructions (it is rolled and out of line:
    // every SASS row of the decode kernel competes for the instruction cache) and still drops the instruction-cache hit
    // rate to 89 %: 25.4 M frames/s against 31.8 M for the three launches.  So the one-kernel path is opt-in (PCB_MC_FUSED=1).
    const bool unfused = getenv("PCB_MC_FUSED") == nullptr || getenv("PCB_MC_UNFUSED") != nullptr;
    const bool fused = h->fast8 && h->quad && !unfused;
    // the tuned list kernel wants its input range-checked: the generator does that on the way (no pre-pass)
    const bool precheck = h->fast8 && h->quad;
    if (precheck) CK(ensure_redo(h, batch));
    const int MC_REDO_CAP = 4096;
    std::vector<cudaEvent_t> ev;
    auto cleanup = [&]() { for (auto e : ev) cudaEventDestroy(e); ev.clear(); };
    auto drain = [&]() -> int {
        CK(cudaStreamSynchronize(h->sc));
        for (size_t i = 0; i < ev.size(); i += 2) { float ms = 0; CK(cudaEventElapsedTime(&ms, ev[i], ev[i + 1])); secs += ms * 1e-3; }
        cleanup();
        return PCB_OK;
    };
    if (fused) {
        const int fpw = scl8q_frames_per_warp(h->L);
        if (!h->d_yscr) CK(cudaMalloc(&h->d_yscr, (size_t)fast_blocks(h) * fpw * h->n * sizeof(float)));
        const size_t spb = scl8q_scratch_per_block(h->m, h->L);
        if (spb && !h->d_qscratch) CK(cudaMalloc(&h->d_qscratch, spb * (size_t)fast_blocks(h)));
        if (!h->d_mc_redo) CK(cudaMalloc(&h->d_mc_redo, (size_t)(MC_REDO_CAP + 1) * sizeof(unsigned long long)));
        CK(cudaMemsetAsync(h->d_mc_redo, 0, sizeof(unsigned long long), h->sc));
        Scl8qGen g;
        g.seed = seed; g.sigma = (float)mc_sigma; g.gcrc = h->d_gcrc; g.gcrc_c0 = h->gcrc_c0;
        g.mc_redo = h->d_mc_redo; g.mc_redo_cap = MC_REDO_CAP;
        CK(ensure_xhat(h, batch));
        if (h->skip_cap < (batch + 31) / 32) {
            cudaFree(h->d_skip); h->d_skip = nullptr; h->skip_cap = 0;
            CK(cudaMalloc(&h->d_skip, ((batch + 31) / 32) * sizeof(uint32_t)));
            h->skip_cap = (batch + 31) / 32;
        }
        g.skip = h->d_skip;
        const char* fenv = getenv("PCB_TEST_FORCE_REDO_MOD");       // test hook: these frame ids take the redo path
        g.force_redo_mod = fenv ? atoi(fenv) : 0;
        for (uint64_t f0 = 0; f0 < frames; f0 += batch) {
            const int nb = (int)(frames - f0 < batch ? frames - f0 : batch);
            g.first_frame = first_frame_id + f0;
            DecodeArgs a = make_args(h, nullptr, 1, nullptr, nullptr, nb);
            CK(cudaMemsetAsync(h->d_redo, 0, 2 * sizeof(int), h->sc));    // [1] is the work counter of the launch
            cudaEvent_t e0, e1;
            if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) { cleanup(); return fail(PCB_ECUDA, "cudaEventCreate failed"); }
            ev.push_back(e0); ev.push_back(e1);
            const int nq = (nb + fpw - 1) / fpw;
            CK(cudaEventRecord(e0, h->sc));
            CK(cudaMemsetAsync(h->d_skip, 0, (((size_t)nb + 31) / 32) * sizeof(uint32_t), h->sc));
            CK(launch_decode_scl8q(a, h->L, h->h_info_w.data(), h->d_xhat, h->d_qscratch, h->d_redo + 1, nullptr, 0, h->d_yscr, 0, &g,
                                   nq < fast_blocks(h) ? nq : fast_blocks(h), h->sc));
            SelectArgs sa = select_args(h, a);
            sa.count = 1; sa.skip = h->d_skip; sa.counters = h->d_counters; sa.seed = seed; sa.first_frame = g.first_frame;
            CK(launch_scl_select(sa, h->sm_count, h->sc));
            CK(cudaEventRecord(e1, h->sc));
            h->launches += 2;
            if (ev.size() >= 512) { if (int r = drain()) return r; }
        }
        if (int r = drain()) return r;
        // frames outside the safe input range of the tuned kernel's arithmetic (|y| tiny or huge: ~1e-9 per frame):
        // the three-launch path decodes them one by one, the generic kernel behind the tuned one
        unsigned long long nredo = 0;
        CK(cudaMemcpy(&nredo, h->d_mc_redo, sizeof nredo, cudaMemcpyDeviceToHost));
        if (nredo > (unsigned long long)MC_REDO_CAP)
            return fail(PCB_ECUDA, "pcb_montecarlo: %llu frames outside the tuned kernel's input range (limit %d per call)", nredo, MC_REDO_CAP);
        std::vector<unsigned long long> ids((size_t)nredo);
        if (nredo) CK(cudaMemcpy(ids.data(), h->d_mc_redo + 1, (size_t)nredo * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        if (nredo) { if (int r = ensure_mc(h, 1)) return r; }
        for (unsigned long long id : ids) {
            GenArgs gg = gen_args(h, snr_db, seed, id, 1);
            CK(launch_generate(gg, h->sm_count, h->sc));
            DecodeArgs a = make_args(h, h->d_ymc, 1, nullptr, h->d_dec, 1);
            CK(launch(h, a, h->sc));
            CK(launch_count(h->d_dec, h->d_tx, h->EW, nbits, 1, h->d_counters, h->sc));
            h->launches += 2;
        }
        CK(cudaStreamSynchronize(h->sc));
    } else {
    if (int r = ensure_mc(h, batch)) return r;
    // everything of a batch is enqueued on one stream and nothing waits for the host between batches; the
    // decode-kernel time comes from one pair of events per batch, read at the end
    for (uint64_t f0 = 0; f0 < frames; f0 += batch) {
        const int nb = (int)(frames - f0 < batch ? frames - f0 : batch);
        GenArgs g = gen_args(h, snr_db, seed, first_frame_id + f0, nb);
        if (precheck) {
            CK(cudaMemsetAsync(h->d_redo, 0, 2 * sizeof(int), h->sc));
            g.redo_count = h->d_redo; g.redo_list = h->d_redo + 2;
        }
        CK(launch_generate(g, h->sm_count, h->sc));
        DecodeArgs a = make_args(h, h->d_ymc, 1, nullptr, h->d_dec, nb);
        a.sigma = mc_sigma;
        cudaEvent_t e0, e1;
        if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) { cleanup(); return fail(PCB_ECUDA, "cudaEventCreate failed"); }
        ev.push_back(e0); ev.push_back(e1);
        CK(cudaEventRecord(e0, h->sc));
        CK(launch(h, a, h->sc, precheck));
        CK(cudaEventRecord(e1, h->sc));
        CK(launch_count(h->d_dec, h->d_tx, h->EW, nbits, nb, h->d_counters, h->sc));
        h->launches += 2;
        if (ev.size() >= 512) { if (int r = drain()) return r; }
    }
    if (int r = drain()) return r;
    }
    unsigned long long cnt[2];
    CK(cudaMemcpy(cnt, h->d_counters, 16, cudaMemcpyDeviceToHost));
    if (frames_done) *frames_done += frames;
    if (frame_errors) *frame_errors += cnt[0];
    if (bit_errors) *bit_errors += cnt[1];
    if (seconds) *seconds += secs;
    return PCB_OK;
}

int pcb_generate_batch_dev(pcb_handle* h, double snr_db, uint64_t seed, uint64_t first_frame_id, int B,
                           float* y_dev, uint32_t* tx_bits_dev) {
    if (!h || !y_dev) return fail(PCB_EINVAL, "pcb_generate_batch_dev: null argument");
    if (!h->code_set) return fail(PCB_ESTATE, "pcb_generate_batch_dev: call pcb_set_code first");
    if (int r = check_gen(h, "pcb_generate_batch_dev")) return r;
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "B < 0");
    if (int r = ensure_device(h)) return r;
    GenArgs g = gen_args(h, snr_db, seed, first_frame_id, B);
    g.y = y_dev; g.tx_bits = tx_bits_dev;
    CK(after_callers_stream(h));
    CK(launch_generate(g, h->sm_count, h->sc));
    h->launches++;
    CK(cudaStreamSynchronize(h->sc));
    return PCB_OK;
}

int pcb_generate_batch(pcb_handle* h, double snr_db, uint64_t seed, uint64_t first_frame_id, int B, float* y,
                       int* tx_bits) {
    if (!h || !y) return fail(PCB_EINVAL, "pcb_generate_batch: null argument");
    if (!h->code_set) return fail(PCB_ESTATE, "pcb_generate_batch: call pcb_set_code first");
    if (int r = check_gen(h, "pcb_generate_batch")) return r;
    if (B <= 0) return B == 0 ? PCB_OK : fail(PCB_EINVAL, "B < 0");
    if (int r = ensure_device(h)) return r;
    if (int r = ensure_mc(h, (size_t)B)) return r;
    GenArgs g = gen_args(h, snr_db, seed, first_frame_id, B);
    CK(launch_generate(g, h->sm_count, h->sc));
    h->launches++;
    CK(cudaMemcpyAsync(y, h->d_ymc, (size_t)B * h->n * 4, cudaMemcpyDeviceToHost, h->sc));
    std::vector<uint32_t> tx((size_t)B * h->EW);
    CK(cudaMemcpyAsync(tx.data(), h->d_tx, tx.size() * 4, cudaMemcpyDeviceToHost, h->sc));
    CK(cudaStreamSynchronize(h->sc));
    if (tx_bits)
        for (int f = 0; f < B; f++)
            for (int t = 0; t < h->stride; t++)
                tx_bits[(size_t)f * h->stride + t] = (tx[(size_t)f * h->EW + (t >> 5)] >> (t & 31)) & 1;
    return PCB_OK;
}


// ---- error counting and the counter reduction (the path's only collective) --------------------
int pcb_count_errors_dev(pcb_handle* h, const uint32_t* dec_bits_dev, const uint32_t* tx_bits_dev, int B, int allreduce,
                         uint64_t* counters) {
    if (!h || !dec_bits_dev || !tx_bits_dev || !counters) return fail(PCB_EINVAL, "pcb_count_errors_dev: null argument");
    if (B < 0) return fail(PCB_EINVAL, "pcb_count_errors_dev: B < 0");
    if (allreduce && !h->comm) return fail(PCB_ESTATE, "pcb_count_errors_dev: allreduce without pcb_comm_init");
    if (int r = ensure_device(h)) return r;
    if (!h->d_red) CK(cudaMalloc(&h->d_red, 8 * sizeof(unsigned long long)));
    const unsigned long long init[3] = {(unsigned long long)B, 0ull, 0ull};
    CK(after_callers_stream(h));
    CK(cudaMemcpyAsync(h->d_red, init, sizeof init, cudaMemcpyHostToDevice, h->sc));
    if (B) {
        CK(launch_count(dec_bits_dev, tx_bits_dev, h->EW, h->decoder == 3 ? h->k : h->n, B, h->d_red + 1, h->sc));
        h->launches++;
    }
    if (allreduce) {
        NcclApi& nc = nccl_api();
        const int rc = nc.AllReduce(h->d_red, h->d_red, 3, NcclApi::kUint64, NcclApi::kSum, h->comm, h->sc);
        if (rc) return fail(PCB_ECUDA, "ncclAllReduce: %s", nc.GetErrorString(rc));
        h->collectives++;
    }
    unsigned long long out[3];
    CK(cudaMemcpyAsync(out, h->d_red, sizeof out, cudaMemcpyDeviceToHost, h->sc));
    CK(cudaStreamSynchronize(h->sc));
    for (int i = 0; i < 3; i++) counters[i] += out[i];
    return PCB_OK;
}

int pcb_comm_unique_id(void* id128) {
    if (!id128) return fail(PCB_EINVAL, "pcb_comm_unique_id: null argument");
    NcclApi& nc = nccl_api();
    if (const char* why = nc.load()) return fail(PCB_ESTATE, "NCCL unavailable: %s", why);
    NcclApi::UniqueId id;
    const int rc = nc.GetUniqueId(&id);
    if (rc) return fail(PCB_ECUDA, "ncclGetUniqueId: %s", nc.GetErrorString(rc));
    memcpy(id128, &id, sizeof id);
    return PCB_OK;
}

int pcb_comm_init(pcb_handle* h, const void* id128, int rank, int world) {
    if (!h || !id128) return fail(PCB_EINVAL, "pcb_comm_init: null argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(PCB_EINVAL, "pcb_comm_init: rank %d of %d", rank, world);
    if (h->comm) return fail(PCB_ESTATE, "pcb_comm_init: the handle already has a communicator");
    if (int r = ensure_device(h)) return r;
    NcclApi& nc = nccl_api();
    if (const char* why = nc.load()) return fail(PCB_ESTATE, "NCCL unavailable: %s", why);
    NcclApi::UniqueId id;
    memcpy(&id, id128, sizeof id);
    const int rc = nc.CommInitRank(&h->comm, world, id, rank);
    if (rc) { h->comm = nullptr; return fail(PCB_ECUDA, "ncclCommInitRank: %s", nc.GetErrorString(rc)); }
    h->comm_rank = rank; h->comm_world = world;
    if (!h->d_red) CK(cudaMalloc(&h->d_red, 8 * sizeof(unsigned long long)));
    return PCB_OK;
}

int pcb_comm_allreduce(pcb_handle* h, uint64_t* counters, int count) {
    if (!h || !counters || count < 1 || count > 8) return fail(PCB_EINVAL, "pcb_comm_allreduce: bad argument");
    if (!h->comm) return fail(PCB_ESTATE, "pcb_comm_allreduce: call pcb_comm_init first");
    if (int r = ensure_device(h)) return r;
    NcclApi& nc = nccl_api();
    CK(cudaMemcpyAsync(h->d_red, counters, count * sizeof(uint64_t), cudaMemcpyHostToDevice, h->sc));
    const int rc = nc.AllReduce(h->d_red, h->d_red, (size_t)count, NcclApi::kUint64, NcclApi::kSum, h->comm, h->sc);
    if (rc) return fail(PCB_ECUDA, "ncclAllReduce: %s", nc.GetErrorString(rc));
    h->collectives++;
    CK(cudaMemcpyAsync(counters, h->d_red, count * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->sc));
    CK(cudaStreamSynchronize(h->sc));
    return PCB_OK;
}


// What bounds pcb_decode_batch (host doubles in): the pinned H2D rate W of this GPU's link and the rate R at
// which this handle's thread pool narrows doubles to floats, measured on the caller's own buffer with the chunk
// size and staging buffers of the pipeline (call it after a pcb_decode_batch; ranks that share a host call it
// together to see the contended figures).  A chunk whose head share x is narrowed on the host costs
// max(x / R, (4x + 8(1 - x)) / W) seconds per value; the bound is that at the best x.
int pcb_host_probe(pcb_handle* h, const double* y_host, int frames, double* h2d_GBps, double* narrow_Gvalues_per_s,
                   double* bound_frames_per_s) {
    if (!h || !y_host || frames < 1) return fail(PCB_EINVAL, "pcb_host_probe: bad argument");
    if (int r = ensure_device(h)) return r;
    if (!h->slot_frames || !h->stage_frames || !h->narrow) return fail(PCB_ESTATE, "pcb_host_probe: run pcb_decode_batch (double input, float decoder) first");
    const size_t chunk = std::min<size_t>({h->slot_frames, h->stage_frames, (size_t)frames});
    const size_t vals = chunk * (size_t)h->n;
    const int reps = (int)std::max<size_t>(1, std::min<size_t>(32, (size_t)frames / chunk));
    h->narrow->start();
    // link: back-to-back copies of distinct chunks, CUDA events around all of them
    CK(cudaEventRecord(h->ev_c0[0], h->sin));
    for (int i = 0; i < reps; i++)
        CK(cudaMemcpyAsync(h->d_y[i & 1], y_host + (size_t)i * vals, vals * 8, cudaMemcpyHostToDevice, h->sin));
    CK(cudaEventRecord(h->ev_c1[0], h->sin));
    // cores: at the same time, as in the pipeline
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < reps; i++) h->narrow->run(y_host + (size_t)i * vals, h->h_stage[i & 1], vals);
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    CK(cudaStreamSynchronize(h->sin));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, h->ev_c0[0], h->ev_c1[0]));
    const double W = (double)reps * vals * 8 / (ms * 1e-3), R = (double)reps * vals / dt;
    double x = 8.0 * R / (W + 4.0 * R);
    x = x > 1.0 ? 1.0 : x;
    const double per_value = std::max(x / R, (4.0 * x + 8.0 * (1.0 - x)) / W);
    if (h2d_GBps) *h2d_GBps = W / 1e9;
    if (narrow_Gvalues_per_s) *narrow_Gvalues_per_s = R / 1e9;
    if (bound_frames_per_s) *bound_frames_per_s = 1.0 / (per_value * h->n);
    return PCB_OK;
}

}  // extern "C"
