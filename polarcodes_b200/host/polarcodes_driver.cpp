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
//   PCB_SEED=<s>     Philox seed (default 0x5EED0000 + SNR index as SURVEY.md §8(d))
//   PCB_CSV=<file>   extra CSV with BER, frames/s, info Gbit/s per SNR point
//   PCB_CONF=<file>  configuration file (default conf.txt in the CWD, like the reference)
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
#include <thread>
#include <vector>

#include "../../include/polar_b200.h"

struct Point {
    uint64_t frames = 0, frame_errors = 0, bit_errors = 0;
    double seconds = 0;
};

static void fail(const char* what) {
    std::fprintf(stderr, "polarcodes_driver: %s: %s\n", what, pcb_last_error());
    std::exit(2);
}

int main() {
    int testType, n, k, decoder, zComputer, L, crc, loop, nth, qFraction;
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
    const char* seed_env = std::getenv("PCB_SEED");
    const int ecrc = decoder == 3 ? crc : 0;   // only the list decoder appends/uses the CRC (:136-155,:442-472)

    std::vector<pcb_handle*> hs(gpus, nullptr);
    for (int g = 0; g < gpus; g++) {
        hs[g] = pcb_create(g, n, k, ecrc, L, decoder, PCB_PREC_STRICT);
        if (!hs[g]) fail("pcb_create");
    }

    std::vector<std::pair<double, double>> res;
    std::vector<Point> points(nth);
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
        std::vector<Point> part(gpus);
        std::vector<std::thread> th;
        for (int g = 0; g < gpus; g++) {
            th.emplace_back([&, g]() {
                if (pcb_set_code(hs[g], mask.data(), sigma) != PCB_OK) fail("pcb_set_code");
                const uint64_t lo = total * g / gpus, hi = total * (g + 1) / gpus;   // shard by global frame id
                if (hi > lo &&
                    pcb_montecarlo(hs[g], snr[si], hi - lo, seed, lo, &part[g].frames, &part[g].frame_errors,
                                   &part[g].bit_errors, &part[g].seconds) != PCB_OK)
                    fail("pcb_montecarlo");
            });
        }
        for (auto& t : th) t.join();
        Point& p = points[si];
        for (int g = 0; g < gpus; g++) {
            p.frames += part[g].frames; p.frame_errors += part[g].frame_errors; p.bit_errors += part[g].bit_errors;
            p.seconds = std::max(p.seconds, part[g].seconds);
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
    std::sort(res.begin(), res.end());
    std::ofstream fout("SC_" + std::to_string(testType) + "_" + std::to_string(n) + "_" + std::to_string(k) + "_L" +
                       std::to_string(L) + "_CRC" + std::to_string(crc) + "_" + std::to_string(time(0)) + ".txt");
    fout << (long long)(endT - startT) << '\n';
    for (size_t i = 0; i < res.size(); i++) fout << res[i].first << '\t' << res[i].second << '\n';

    if (const char* csv = std::getenv("PCB_CSV")) {
        std::ofstream c(csv);
        c << "snr_db,frames,frame_errors,bit_errors,fer,ber,decode_seconds,frames_per_s,info_gbit_per_s\n";
        const double bits = decoder == 3 ? k : n;
        for (int si = 0; si < nth; si++) {
            const Point& p = points[si];
            const double fps = p.seconds > 0 ? p.frames / p.seconds : 0;
            c << snr[si] << ',' << p.frames << ',' << p.frame_errors << ',' << p.bit_errors << ','
              << (p.frames ? (double)p.frame_errors / p.frames : 0) << ','
              << (p.frames ? (double)p.bit_errors / (p.frames * bits) : 0) << ',' << p.seconds << ',' << fps << ','
              << fps * k / 1e9 << '\n';
        }
    }
    for (auto h : hs) pcb_destroy(h);
    return 0;
}
