// polarcodes_driver — batched conf.txt-driven Monte-Carlo loop on top of the C ABI.
//
// Drop-in for the reference's main()/myThread() (PolarCodes.cpp:94-308): reads the same
// conf.txt (10 ints then the SNR list, whitespace separated, same order, :257-284), echoes the
// same prompts, and writes the same result file `SC_<type>_<n>_<k>_L<L>_CRC<crc>_<time>.txt`
// (first line elapsed clock ticks, then `snr \t value` per point, :294-307).  The per-frame loop
// body (:132-192) is replaced by pcb_montecarlo: frames are generated (Philox BPSK/AWGN),
// encoded, decoded and compared on the GPU in batches.
//
// Additions that keep an unmodified conf.txt working (all optional, via environment):
//   PCB_GPUS=<g>     use GPUs 0..g-1 of this box (one host thread + handle per GPU); frames are
//                    sharded by global frame id, counters summed on the host (24 bytes per SNR
//                    point; bench.py does the same reduction with NCCL across processes)
//   one process per GPU (torchrun --no-python, mpirun, or a shell loop): with WORLD_SIZE > 1 in the
//                    environment (RANK, LOCAL_RANK as torchrun sets them) every process drives GPU
//                    LOCAL_RANK, decodes its share of each slice and the ranks sum
//                    {frames, frame errors, bit errors} with ncclAllReduce after every slice — that
//                    is also the poll of the cross-rank error budget (PCB_MAX_ERRORS).  Rank 0 hands
//                    the NCCL unique id to the others through the file PCB_NCCL_ID_FILE (default
//                    /tmp/pcb_nccl_id.<parent pid>.<MASTER_PORT>), prints, checkpoints and writes the
//                    result file; the counters equal those of a one-GPU run (Philox is keyed by the
//                    global frame id).
//   PCB_PREC=fp32|int8  decoder 2 only: the fp32 fast mode, or the 8-bit mode with the scale 2 << qFraction of conf.txt line 9
//                    (the reference reads qFraction and never uses it: SCLDecoderMinsum.cpp:405); default = the reference's arithmetic
//   PCB_SEED=<s>     Philox seed (default 0x5EED0000 + SNR index as SURVEY.md §8(d))
//   PCB_CSV=<file>   extra CSV with BER, frames/s, info Gbit/s per SNR point
//   PCB_CONF=<file>  configuration file (default conf.txt in the CWD, like the reference)
//   PCB_CKPT=<file>  checkpoint of a long run: after every slice of
//                    PCB_SLICE frames (default 2^22) the per-point counters and the frame-id
//                    cursor are written (tmp + rename); a later run with the same configuration
//                    resumes behind the cursor.  The Philox stream is keyed by the global frame
//                    id, so a resumed run ends with exactly the counters of an uninterrupted one.
//   PCB_MAX_ERRORS=<e>  error budget: an SNR point stops at the first slice boundary with at
//                    least e frame errors (FER = errors / frames decoded so far)
//   PCB_STOP_AFTER_SLICES=<s>  exit with status 3 after s slices (interruption drill for the tests)
// testType 2 (throughput) follows the reference semantics of timing the decoder only: the value
// written is the decode-kernel time in clock ticks (CLOCKS_PER_SEC) accumulated over SNR points
// (the reference never resets tAll, :118,243), for `loop` distinct frames.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <iostream>
#include <string>
#include <chrono>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <vector>

#include "../../include/polar_b200.h"

struct Point {
    double lane_ops = 0;   // algorithmic ALU lane-ops per frame of this point's mask (not checkpointed: recomputed)
    uint64_t frames = 0, frame_errors = 0, bit_errors = 0;
    double seconds = 0;
    uint64_t cursor = 0;   // global frame ids below the cursor are accounted for
    int done = 0;
};

// Checkpoint: one header line that identifies the run, then one line per SNR point.
static std::string ckpt_header(int testType, int n, int k, int loop, int decoder, int zComputer, int L, int crc,
                               const std::vector<double>& snr, const char* seed_env) {
    char buf[256];
    std::snprintf(buf, sizeof buf, "pcb_ckpt 1 %d %d %d %d %d %d %d %d %s %zu", testType, n, k, loop, decoder, zComputer, L, crc,
                  seed_env ? seed_env : "default", snr.size());
    std::string h = buf;
    for (double v : snr) { std::snprintf(buf, sizeof buf, " %.17g", v); h += buf; }
    return h;
}

static bool ckpt_load(const char* path, const std::string& header, std::vector<Point>& pts) {
    std::ifstream f(path);
    if (!f) return false;
    std::string h;
    std::getline(f, h);
    if (h != header) {
        std::fprintf(stderr, "polarcodes_driver: checkpoint %s belongs to another configuration, ignored\n", path);
        return false;
    }
    std::vector<Point> tmp(pts.size());
    for (auto& p : tmp)
        if (!(f >> p.cursor >> p.frames >> p.frame_errors >> p.bit_errors >> p.seconds >> p.done)) return false;
    pts = tmp;
    return true;
}

static void ckpt_save(const char* path, const std::string& header, const std::vector<Point>& pts) {
    const std::string tmp = std::string(path) + ".tmp";
    {
        std::ofstream f(tmp);
        f << header << '\n';
        for (const auto& p : pts) {
            char buf[200];
            std::snprintf(buf, sizeof buf, "%llu %llu %llu %llu %.17g %d\n", (unsigned long long)p.cursor, (unsigned long long)p.frames,
                          (unsigned long long)p.frame_errors, (unsigned long long)p.bit_errors, p.seconds, p.done);
            f << buf;
        }
    }
    std::rename(tmp.c_str(), path);
}

static void fail(const char* what) {
    std::fprintf(stderr, "polarcodes_driver: %s: %s\n", what, pcb_last_error());
    std::exit(2);
}

static int env_int(const char* name, int dflt) {
    const char* e = std::getenv(name);
    return e && *e ? std::atoi(e) : dflt;
}

// Rank 0 publishes the NCCL unique id in a file (tmp + rename); the other ranks wait for a fresh one.
static void exchange_id(int rank, char (&id)[128]) {
    std::string path;
    if (const char* e = std::getenv("PCB_NCCL_ID_FILE")) path = e;
    else {
        const char* port = std::getenv("MASTER_PORT");
        path = "/tmp/pcb_nccl_id." + std::to_string((long)getppid()) + "." + (port ? port : "0");
    }
    if (rank == 0) {
        if (pcb_comm_unique_id(id) != PCB_OK) fail("pcb_comm_unique_id");
        const std::string tmp = path + ".tmp";
        { std::ofstream f(tmp, std::ios::binary); f.write(id, 128); }
        std::rename(tmp.c_str(), path.c_str());
        return;
    }
    const time_t started = time(nullptr);
    for (int tries = 0; tries < 6000; tries++) {      // up to ten minutes
        struct stat st;
        if (stat(path.c_str(), &st) == 0 && st.st_size == 128 && st.st_mtime >= started - 30) {
            std::ifstream f(path, std::ios::binary);
            if (f.read(id, 128)) return;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
    std::fprintf(stderr, "polarcodes_driver: rank %d never saw the NCCL id file %s\n", rank, path.c_str());
    std::exit(2);
}

int main() {
    int testType, n, k, decoder, zComputer, L, crc, loop, nth, qFraction;
    const int world = std::max(1, env_int("WORLD_SIZE", 1)), rank = env_int("RANK", 0), local_rank = env_int("LOCAL_RANK", rank);
    if (rank != 0) std::cout.setstate(std::ios::failbit);      // one voice: rank 0 prints, checkpoints and writes the files
    const char* conf = std::getenv("PCB_CONF");
    std::ifstream fin(conf ? conf : "conf.txt");
    if (!fin) { std::fprintf(stderr, "polarcodes_driver: cannot open %s\n", conf ? conf : "conf.txt"); return 2; }
    std::cout << "Input test type FER (1), throughput (2)\n";
    fin >> testType;
    std::cout << testType << '\n';
    std::cout << "Input n, k, number of iteration\n";
    fin >> n >> k >> loop;
    std::cout << n << " " << k << " " << loop << '\n';
    std::cout << "Select decoder: SC (1), SCMinsum (2), SCLMinsum (3)\n";
    fin >> decoder;
    std::cout << decoder << '\n';
    std::cout << "Select zComputer: BEC (1), AWGN (2)\n";
    fin >> zComputer;
    std::cout << zComputer << '\n';
    std::cout << "Input L\n";
    fin >> L;
    std::cout << L << '\n';
    std::cout << "Input CRC\n";
    fin >> crc;
    std::cout << crc << '\n';
    std::cout << "Input Q fraction\n";
    fin >> qFraction;
    std::cout << qFraction << '\n';
    std::cout << "Input number of SNRb\n";
    fin >> nth;
    std::cout << nth << '\n';
    if (!fin || nth < 0) { std::fprintf(stderr, "polarcodes_driver: malformed configuration\n"); return 2; }
    std::vector<double> snr(nth);
    std::cout << "Input " << nth << " SNRbs\n";
    for (int i = 0; i < nth; i++) {
        fin >> snr[i];
        std::cout << snr[i] << " ";
    }
    std::cout << "\nRunning...\n\n";

    int gpus = 1;
    if (const char* g = std::getenv("PCB_GPUS")) gpus = std::max(1, std::atoi(g));
    gpus = std::min(gpus, std::max(1, pcb_device_count()));
    if (world > 1) gpus = 1;                                   // one process per GPU
    const char* seed_env = std::getenv("PCB_SEED");
    // The loop appends the CRC to the payload and builds InfIndexesAWGN(n, k + crc) for EVERY decoder
    // (:125, :136-155); only the list decoder checks it (:442-472).  The BEC construction ignores crc
    // (InfIndexesBEC(n, k), :125): with crc > 0 the reference then over-reads its frozen vector (:127,
    // Encoder.cpp:17-33) — undefined there; here decoders 1, 2 run the k-position code and decoder 3 is refused.
    const int ecrc = zComputer == 1 && decoder != 3 ? 0 : crc;

    std::vector<pcb_handle*> hs(gpus, nullptr);
    for (int g = 0; g < gpus; g++) {
        int prec = PCB_PREC_STRICT;
        if (const char* pe = std::getenv("PCB_PREC")) {
            if (decoder == 2 && std::string(pe) == "fp32") prec = PCB_PREC_FP32;
            else if (decoder == 2 && std::string(pe) == "int8") prec = PCB_PREC_INT8_Q(qFraction);
            else if (*pe) { std::fprintf(stderr, "PCB_PREC=%s: decoder 2 takes fp32 or int8\n", pe); return 2; }
        }
        hs[g] = pcb_create(world > 1 ? local_rank : g, n, k, ecrc, L, decoder, prec);
        if (!hs[g]) fail("pcb_create");
    }
    if (world > 1) {
        char id[128];
        exchange_id(rank, id);
        if (pcb_comm_init(hs[0], id, rank, world) != PCB_OK) fail("pcb_comm_init");
    }

    std::vector<std::pair<double, double>> res;
    std::vector<Point> points(nth);
    const char* ckpt = std::getenv("PCB_CKPT");
    const std::string header = ckpt_header(testType, n, k, loop, decoder, zComputer, L, crc, snr, seed_env);
    if (ckpt && ckpt_load(ckpt, header, points)) std::cout << "Resuming from " << ckpt << "\n";   // every rank reads the same file
    uint64_t slice = 1ull << 22;
    if (const char* e = std::getenv("PCB_SLICE")) slice = std::max<uint64_t>(1, std::strtoull(e, nullptr, 0));
    uint64_t max_errors = 0, stop_after = 0, slices_done = 0;
    if (const char* e = std::getenv("PCB_MAX_ERRORS")) max_errors = std::strtoull(e, nullptr, 0);
    if (const char* e = std::getenv("PCB_STOP_AFTER_SLICES")) stop_after = std::strtoull(e, nullptr, 0);
    const clock_t startT = clock();
    double tAll = 0;
    for (int si = 0; si < nth; si++) {
        // code construction per SNR point, design SNR = operating SNR (:123-128)
        const double sigma = pcb_sigma(n, k, snr[si]);
        const int ninf = zComputer == 1 ? k : k + ecrc;
        std::vector<int> inf(ninf), mask(n);
        if ((zComputer == 1 ? pcb_inf_indexes_bec(n, k, inf.data()) : pcb_inf_indexes_awgn(n, ninf, sigma, inf.data())) != PCB_OK)
            fail("code construction");
        if (pcb_frozen_mask(n, inf.data(), ninf, mask.data()) != PCB_OK) fail("pcb_frozen_mask");
        const uint64_t seed = seed_env ? std::strtoull(seed_env, nullptr, 0) : (0x5EED0000ull + (uint64_t)si);
        const uint64_t total = (uint64_t)std::max(loop, 0);
        Point& p = points[si];
        if (pcb_lane_ops_per_frame(mask.data(), n, L, decoder, &p.lane_ops, nullptr) != PCB_OK) fail("pcb_lane_ops_per_frame");
        bool code_set = false;
        while (!p.done) {
            if (p.cursor >= total || (max_errors && p.frame_errors >= max_errors)) {
                p.done = 1;
                if (ckpt && rank == 0) ckpt_save(ckpt, header, points);
                break;
            }
            if (stop_after && slices_done >= stop_after) {
                std::cout << "\nStopped after " << slices_done << " slices\n";
                return 3;
            }
            // one slice of global frame ids [base, base + len), sharded over the GPUs
            const uint64_t base = p.cursor, len = std::min(slice, total - base);
            std::vector<Point> part(gpus);
            std::vector<std::thread> th;
            for (int g = 0; g < gpus; g++) {
                th.emplace_back([&, g]() {
                    if (!code_set && pcb_set_code(hs[g], mask.data(), sigma) != PCB_OK) fail("pcb_set_code");
                    const int part_i = world > 1 ? rank : g, parts = world > 1 ? world : gpus;
                    const uint64_t lo = base + len * part_i / parts, hi = base + len * (part_i + 1) / parts;
                    if (hi > lo &&
                        pcb_montecarlo(hs[g], snr[si], hi - lo, seed, lo, &part[g].frames, &part[g].frame_errors,
                                       &part[g].bit_errors, &part[g].seconds) != PCB_OK)
                        fail("pcb_montecarlo");
                });
            }
            for (auto& t : th) t.join();
            code_set = true;
            double sec = 0;
            if (world > 1) {
                // the reference's `er` counter (:129,165,174,183,193) summed over the ranks; every rank gets the
                // totals, so every rank takes the same stop decision at the next slice boundary
                uint64_t c[4] = {part[0].frames, part[0].frame_errors, part[0].bit_errors, (uint64_t)(part[0].seconds * 1e9)};
                if (pcb_comm_allreduce(hs[0], c, 4) != PCB_OK) fail("pcb_comm_allreduce");
                p.frames += c[0]; p.frame_errors += c[1]; p.bit_errors += c[2];
                sec = (double)c[3] * 1e-9 / world;              // mean decode-kernel time of the ranks
            } else {
                for (int g = 0; g < gpus; g++) {
                    p.frames += part[g].frames; p.frame_errors += part[g].frame_errors; p.bit_errors += part[g].bit_errors;
                    sec = std::max(sec, part[g].seconds);
                }
            }
            p.seconds += sec;
            p.cursor = base + len;
            slices_done++;
            if (ckpt && rank == 0) ckpt_save(ckpt, header, points);
        }
        if (testType == 1) {
            const double fer = p.frames ? (double)p.frame_errors / (double)p.frames : 0.0;
            res.push_back({snr[si], fer});
            std::cout << '\n' << snr[si] << " " << fer << '\n';
        } else {
            tAll += p.seconds * CLOCKS_PER_SEC;
            res.push_back({snr[si], (double)(long long)tAll});
            std::cout << '\n' << snr[si] << " " << (long long)tAll << '\n';
        }
    }
    const clock_t endT = clock();
    if (rank != 0) {                                           // rank 0 alone writes the result file and the CSV
        for (auto h : hs) pcb_destroy(h);
        return 0;
    }
    std::sort(res.begin(), res.end());
    std::ofstream fout("SC_" + std::to_string(testType) + "_" + std::to_string(n) + "_" + std::to_string(k) + "_L" +
                       std::to_string(L) + "_CRC" + std::to_string(crc) + "_" + std::to_string(time(0)) + ".txt");
    fout << (long long)(endT - startT) << '\n';
    for (size_t i = 0; i < res.size(); i++) fout << res[i].first << '\t' << res[i].second << '\n';

    if (const char* csv = std::getenv("PCB_CSV")) {
        std::ofstream c(csv);
        c << "snr_db,frames,frame_errors,bit_errors,fer,ber,decode_seconds,frames_per_s,info_gbit_per_s,lane_ops_per_frame,alu_roofline_frac\n";
        const double bits = decoder == 3 ? k : n;
        double peak = 0;                                        // SMs x lanes x max SM clock of GPU 0, times the GPUs
        if (pcb_alu_peak(hs[0], &peak) != PCB_OK) fail("pcb_alu_peak");
        peak *= world > 1 ? world : gpus;
        for (int si = 0; si < nth; si++) {
            const Point& p = points[si];
            const double fps = p.seconds > 0 ? p.frames / p.seconds : 0;
            c << snr[si] << ',' << p.frames << ',' << p.frame_errors << ',' << p.bit_errors << ','
              << (p.frames ? (double)p.frame_errors / p.frames : 0) << ','
              << (p.frames ? (double)p.bit_errors / (p.frames * bits) : 0) << ',' << p.seconds << ',' << fps << ','
              << fps * k / 1e9 << ',' << p.lane_ops << ',' << (peak > 0 ? fps * p.lane_ops / peak : 0) << '\n';
        }
    }
    for (auto h : hs) pcb_destroy(h);
    return 0;
}
