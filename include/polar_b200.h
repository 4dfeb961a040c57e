/* polar_b200.h — C ABI of the B200-native batched polar encoder / SC / SCL decoder.
 *
 * This is the drop-in boundary for the hot path of dim49v/polarCodes (the per-frame loop body
 * of PolarCodes.cpp:132-192).  The reference has no FFI layer: its "plugin API" is the set of
 * free functions the Monte-Carlo loop calls.  Each entry point below cites the reference
 * function it replaces (paths relative to /root/reference/PolarCodes/).  Plain pointers and
 * sizes only; every pointer is a HOST pointer unless the name ends in _dev.
 *
 * Conventions: functions return 0 on success, a negative PCB_E* code on failure, and
 * pcb_last_error() returns a thread-local human readable message.  A handle owns all device
 * memory and streams for one (GPU, code configuration); calls on one handle must be
 * serialised by the caller (same rule as the reference's namespace-global decoder state,
 * SCLDecoderMinsum.cpp:17-38).  There is NO CPU fallback: without a CUDA device pcb_create
 * fails.
 */
#ifndef POLAR_B200_H
#define POLAR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCB_OK 0
#define PCB_EINVAL (-1)   /* bad argument / unsupported configuration */
#define PCB_ECUDA (-2)    /* CUDA runtime error (message in pcb_last_error) */
#define PCB_ENODEV (-3)   /* no CUDA device */
#define PCB_ESTATE (-4)   /* call order (e.g. decode before pcb_set_code) */

/* decoder ids are the reference's conf.txt values (PolarCodes.cpp:159-188) */
#define PCB_DEC_SC 1        /* SCDecoder::Decode        SCDecoder.cpp:74        fp64 tanh/atanh */
#define PCB_DEC_SCMINSUM 2  /* SCDecoderMinsum::Decode  SCDecoderMinsum.cpp:68  fp64 min-sum   */
#define PCB_DEC_SCLMINSUM 3 /* SCLDecoderMinsum::Decode SCLDecoderMinsum.cpp:401 fp32 list     */

/* precision: 0 = strict (the reference's arithmetic type: fp64 for decoders 1,2, fp32 for 3),
 *            1 = fp32 fast mode for decoder 2 (not bit-exact by contract; FER checked statistically) */
#define PCB_PREC_STRICT 0
#define PCB_PREC_FP32 1
/* 2 = 8-bit min-sum for decoder 2, four LLRs per 32-bit register: the mode the reference sketches and never wires up
 * (Decodef_char / Decodeg_char, SCDecoderMinsum.cpp:17-30; the scale `2 << qFraction`, SCLDecoderMinsum.cpp:405,
 * conf.txt:9).  q(y) = clamp(rint(y * (2 << q)), -127, 127), f = Decodef_char, g = Decodeg_char with the sum clamped to
 * +-127 instead of wrapping.  A mode of its own, with its own FER curve (not bit-exact with the double decoder; bit-exact
 * with its definition, po_sc8_decode_batch of the CPU restatement under oracle/).  n in [64, 4096].  Pass PCB_PREC_INT8_Q(qFraction). */
#define PCB_PREC_INT8 2
#define PCB_PREC_INT8_Q(q) (PCB_PREC_INT8 | ((q) << 8))

typedef struct pcb_handle pcb_handle;

const char* pcb_last_error(void);
const char* pcb_version(void);
int pcb_device_count(void);

/* ---- code construction (host side, runs once per SNR point) -------------------------- */
/* PolarCodes.cpp:123  sigma = sqrt(n / (2 k 10^(snr/10))) */
double pcb_sigma(int n, int k, double snr_db);
/* PolarCodes.cpp:59-74 InfIndexesAWGN: the k most reliable positions, ascending. out[k] */
int pcb_inf_indexes_awgn(int n, int k, double sigma, int* out);
/* PolarCodes.cpp:77-92 InfIndexesBEC. out[k] */
int pcb_inf_indexes_bec(int n, int k, int* out);
/* Encoder.cpp:43-55 FrozenBits with the all-zero frozen vector of PolarCodes.cpp:127:
 * mask[i] = -1 for an information position, else the frozen value (0). mask[n] */
int pcb_frozen_mask(int n, const int* inf, int ninf, int* mask);
/* Encoder.cpp:17-33 AddFrozenBits (frozen values all zero). x[n] */
int pcb_add_frozen_bits(const int* xinf, int* x, int n, const int* inf, int ninf);
/* Resources.cpp:108-122 / :124-138 — one int per bit, MSB first, len a multiple of 8 */
unsigned pcb_crc8(const int* bits, int len);
unsigned pcb_crc16(const int* bits, int len);
/* PolarCodes.cpp:136-155 — append the crc (0/8/16) MSB first at xinf[k..k+crc) */
int pcb_append_crc(int* xinf, int k, int crc);

/* Algorithmic ALU lane-ops of one frame (the roofline numerator): counted from the frozen mask with the
 * fixed weights f = 4, g = 2, metric update = 2, compare-exchange = 3 (bitonic convention over 2L
 * candidates), 1 per packed word of partial-sum XORs; decoders 1,2: L = 1, no metric, no select.
 * llr_updates (may be NULL) receives the number of f + g updates. */
int pcb_lane_ops_per_frame(const int* frozen_mask, int n, int L, int decoder, double* lane_ops, double* llr_updates);

/* ---- handle -------------------------------------------------------------------------- */
/* Replaces the per-run allocations of PolarCodes.cpp:99-121 and
 * SCLDecoderMinsum::InitializeDataStructures (SCLDecoderMinsum.cpp:60-118).
 * n: power of two in [2, 65536]; L: power of two in [1, 32] (decoder 3; ignored otherwise);
 * crc in {0, 8, 16}: checked by decoder 3 (:442-472); for decoders 1,2 it only tells the frame
 * generator to append it to the payload, as the loop does for every decoder (PolarCodes.cpp:136-155).
 * k is the payload size: sigma / the rate use k, never k + crc (PolarCodes.cpp:123).
 * device: CUDA ordinal. Returns NULL on failure. */
pcb_handle* pcb_create(int device, int n, int k, int crc, int L, int decoder, int precision);
void pcb_destroy(pcb_handle* h);

/* Uploads the frozen mask (the vector FrozenBits() returns, Encoder.cpp:43) and sigma (used by
 * decoder 1 only: LLR = 2y/sigma^2, SCDecoder.cpp:15-17).  The number of -1 entries must be k + crc
 * for decoder 3; decoders 1,2 decode a mask with any count, but pcb_montecarlo / pcb_generate_*
 * need k + crc positions to scatter the payload into (PCB_ESTATE otherwise). */
int pcb_set_code(pcb_handle* h, const int* frozen_mask, double sigma);

/* Encode (Encoder.cpp:77-83), batched: u[B][n] one int per bit -> x[B][n].
 * with_shuffle != 0: x = bitrev(u) F^(x)m (used with decoders 1,2); 0: x = u F^(x)m (decoder 3). */
int pcb_encode_batch(pcb_handle* h, const int* u, int* x, int with_shuffle, int B);

/* The same on packed frames (the throughput path: n / 4 bytes per frame over the link or HBM instead of 8 n):
 * u_bits[B][words] -> x_bits[B][words], words = max(n / 32, 1), bit t of a frame at word t / 32, bit t % 32.
 * _packed: host buffers, chunked and double-buffered; _dev: device buffers, ms (may be NULL) = kernel time. */
int pcb_encode_batch_packed(pcb_handle* h, const uint32_t* u_bits, uint32_t* x_bits, int with_shuffle, int B);
int pcb_encode_batch_dev(pcb_handle* h, const uint32_t* u_bits_dev, uint32_t* x_bits_dev, int with_shuffle, int B, float* ms);

/* Decode B frames. y[B][n] are the raw channel values the reference passes to *::Decode
 * (PolarCodes.cpp:162,171,180).  sol is [B][n] for decoders 1,2 (full u-hat incl. frozen
 * positions) and [B][k+crc] for decoder 3 (info + crc bits), one int per bit.
 * Host buffers (pinned ones overlap best), chunked H2D / kernel / D2H pipeline.  For the float
 * decoders (2 non-strict, 3) the (float) narrowing the reference applies to y is done on the way:
 * part of each chunk on the host cores (PCB_HOST_THREADS, default = usable CPUs / ranks of the
 * node - 2), the rest on the device, split so that the cores and the PCIe link finish together
 * (PCB_NO_HOST_NARROW=1: every value crosses the link as a double). */
int pcb_decode_batch(pcb_handle* h, const double* y, int* sol, int B);
/* Same with fp32 channel values. Bit-exact with pcb_decode_batch for decoder 3 when
 * yf[i] == (float)y[i] (SCLDecoderMinsum.cpp:406-408 does that conversion itself). */
int pcb_decode_batch_f32(pcb_handle* h, const float* y, int* sol, int B);
/* Packed output variant: sol_bits[B][words] with words = ceil(stride/32), bit t of the frame's
 * solution at word t/32, bit t%32 (stride = n or k+crc as above). */
int pcb_decode_batch_f32_packed(pcb_handle* h, const float* y, uint32_t* sol_bits, int B);
/* Device-resident variant: y_dev is a device pointer ([B][n] float if y_is_f32 else double),
 * sol_bits_dev a device pointer to [B][words] uint32. Runs on the handle's stream and
 * synchronises before returning. ms (may be NULL) receives the kernel time from CUDA events.
 * Stream ordering of every *_dev entry point: the handle's stream is non-blocking; the call is
 * ordered after the work already queued on the legacy default stream (where a torch caller's
 * tensors are produced), and it returns with its own work complete. A caller that fills the
 * buffers on other non-blocking streams synchronises those before calling. */
int pcb_decode_batch_dev(pcb_handle* h, const void* y_dev, int y_is_f32, uint32_t* sol_bits_dev, int B,
                         float* ms);

/* No-allocation variant of pcb_decode_batch_dev.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and the same (bit-identical) output, but the call makes
 * no cudaMalloc / cudaFree / allocating library call: it runs the reference-literal kernel (one warp per
 * frame) with its per-warp LLR scratch in a static device array of PCB_NOALLOC_SCRATCH_BYTES that every GPU
 * which loads the library holds until exit.  Limits: B <= PCB_NOALLOC_MAX_FRAMES per call, and one block's
 * scratch (about 4 n L bytes per warp, 8 n for the fp64 decoders) must fit the array — every code a handle
 * can be created for does today (at most 3.5 MiB), the check guards later changes; beyond either limit the
 * call returns PCB_EINVAL with the numbers in pcb_last_error().  The grid is the number of blocks whose
 * scratch fits the array, so the call is slower than pcb_decode_batch_dev.  All calls on a GPU share the array (whatever the handle): calls that
 * overlap on the GPU race; serialise them.  Synchronises before returning. */
#define PCB_NOALLOC_SCRATCH_BYTES (8u << 20)
#define PCB_NOALLOC_MAX_FRAMES (1 << 20)
int pcb_decode_batch_dev_noalloc(pcb_handle* h, const void* y_dev, int y_is_f32, uint32_t* sol_bits_dev, int B);

/* Batched Monte-Carlo body (PolarCodes.cpp:132-192) with on-device Philox4x32-10 BPSK/AWGN:
 * frames [first_frame_id, first_frame_id + frames) of the stream keyed by `seed`; results do
 * not depend on batch size or GPU count. Counters are ADDED to *frame_errors / *bit_errors
 * (bit errors over the k info bits for decoder 3, over all n u-bits for decoders 1,2 — the
 * reference counts frame errors only, PolarCodes.cpp:163-186). seconds = decode-kernel time.
 * The channel values are fp32 (s + sigma z with z from Philox4x32-10 + Box-Muller in fp32, |z| up to
 * 6.76: a 2^-33 floor on the uniform); the reference draws N(0, sigma) in double from mt19937.  Parity
 * of this mode is statistical by construction; FER points below ~1e-9 per bit would
 * feel the truncated tail.  The handle's own sigma (pcb_set_code) is not touched. */
int pcb_montecarlo(pcb_handle* h, double snr_db, uint64_t frames, uint64_t seed, uint64_t first_frame_id,
                   uint64_t* frames_done, uint64_t* frame_errors, uint64_t* bit_errors, double* seconds);

/* Generate one batch of channel values on the device exactly as pcb_montecarlo would and copy
 * it to the host (verification hook: feed the result to the oracle). y[B][n] float,
 * xinf[B][k+crc] int (decoder 3) or u[B][n] (decoders 1,2). */
int pcb_generate_batch(pcb_handle* h, double snr_db, uint64_t seed, uint64_t first_frame_id, int B, float* y,
                       int* tx_bits);

/* Same, into caller-provided DEVICE buffers: y_dev[B][n] float, tx_bits_dev[B][ceil(stride/32)]
 * packed (may be NULL). Synchronises before returning. */
int pcb_generate_batch_dev(pcb_handle* h, double snr_db, uint64_t seed, uint64_t first_frame_id, int B,
                           float* y_dev, uint32_t* tx_bits_dev);

/* ---- error counters and their reduction across GPUs ------------------------------------------
 * Frame-error compare of the loop (PolarCodes.cpp:163-168,172-177,181-186) on packed DEVICE words:
 * dec_bits_dev / tx_bits_dev are [B][ceil(stride/32)] (what pcb_decode_batch_dev writes and
 * pcb_generate_batch_dev hands out).  counters[3] += {B, frame errors, bit errors} (bit errors over
 * the k info bits for decoder 3, all n u-bits for decoders 1,2).  allreduce != 0: the three numbers
 * are first summed over the ranks of the communicator (ncclAllReduce on the handle's stream). */
int pcb_count_errors_dev(pcb_handle* h, const uint32_t* dec_bits_dev, const uint32_t* tx_bits_dev, int B, int allreduce,
                         uint64_t* counters);
/* One process per GPU: rank 0 obtains a 128-byte NCCL unique id, hands it to the other ranks by any
 * means (a file, MPI, torch.distributed), and every rank binds a communicator to its handle.  The
 * `er` counter of the reference loop (PolarCodes.cpp:129,193) then becomes
 * pcb_comm_allreduce(SUM, uint64[count <= 8]).  NCCL is opened at run time (libnccl.so.2). */
int pcb_comm_unique_id(void* id128);
int pcb_comm_init(pcb_handle* h, const void* id128, int rank, int world);
int pcb_comm_allreduce(pcb_handle* h, uint64_t* counters, int count);

/* ALU roofline of this handle's device for this decoder's arithmetic: SMs x lanes x max SM clock, lanes =
 * 128 per SM (fp32 / int32) or 64 (fp64: decoders 1,2 in strict precision).  lane-ops per second. */
int pcb_alu_peak(pcb_handle* h, double* lane_ops_per_s);

/* The two rates that bound pcb_decode_batch with double host input on a float decoder: the pinned H2D rate of
 * this GPU and the double->float narrowing rate of the handle's host thread pool (y_host: `frames` frames of
 * pinned doubles, read only), and the frames/s they allow when combined as the pipeline combines them.
 * Call after a pcb_decode_batch; ranks that share a host call it together. */
int pcb_host_probe(pcb_handle* h, const double* y_host, int frames, double* h2d_GBps, double* narrow_Gvalues_per_s,
                   double* bound_frames_per_s);

/* introspection for bench / roofline accounting */
int pcb_info(pcb_handle* h, int* sm_count, int* warps_per_block, int* blocks, int* smem_per_block,
             int* first_smem_layer, uint64_t* kernel_launches);

#ifdef __cplusplus
}
#endif
#endif /* POLAR_B200_H */
