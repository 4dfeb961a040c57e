#!/usr/bin/env python
"""bench.py — headline benchmark of the polar decode hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the decoder over one batch of synthetic frames (on-device Philox BPSK/AWGN
codewords of the AWGN-constructed code, 2.0 dB) followed by the error count of that batch and — at
N > 1 — the NCCL all-reduce of the three counters through the product's own communicator
(pcb_count_errors_dev), i.e. the loop body of PolarCodes.cpp:132-193 for a whole batch.  Workload:
SCLDecoderMinsum L=8, N=1024, K=512 (BASELINE.json configs[2], the configuration the metric is quoted on).
  value     decoded info Gbit/s with the channel values (double, the reference's type) resident in HBM
  e2e       same metric through the reference-facing C-ABI call pcb_decode_batch with HOST (pinned) double
            buffers: H2D + kernels + D2H inside the timed region; `bound` = what the measured PCIe and
            host-narrowing rates allow
  roofline  ALU lane-ops (SURVEY.md 8(d) convention, counted exactly from the frozen mask) / kernel time
            from CUDA events on the launch stream / (SMs * 128 lanes * max SM clock)
  configs   the other BASELINE.json configurations (A double / A' fp32, B, C crc16, D, E): frames/s, Gbit/s,
            their own ALU-roofline fraction (fp64 decoders against 64 lanes per SM) and DRAM bytes per
            algorithmic byte from the committed ncu capture — N = 1 only
  montecarlo  the drop-in for the reference loop: pcb_montecarlo (Philox -> CRC -> encode -> decode -> count
            on the device, nothing crosses PCIe), wall clock of the call — N = 1 only
  cpu_baseline  the unmodified reference (oracle/_ref) or the C port timed on the host cores — N = 1 only
Multi-GPU: one process per GPU (torchrun), frames sharded, no data-path collective (weak scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N, K, LIST, SNR_DB = 1024, 512, 8, 2.0
METRIC = "decoded info Gbit/s, SCL-8 N=1024 K=512"


def workload_config(crc):
    """Identical in both arms (ours / --impl reference): the driver compares the two `config` objects."""
    return {"workload": f"SCLDecoderMinsum L={LIST} N={N} K={K} crc={crc} AWGN-construction {SNR_DB} dB",
            "decoder": "SCLDecoderMinsum", "n": N, "k": K, "L": LIST, "crc": crc, "snr_db": SNR_DB,
            "construction": "AWGN (InfIndexesAWGN at the operating SNR)", "input": "double y[B][n], BPSK + AWGN"}


def lane_ops_per_frame(mask, L):
    """f=4, g=2 lane-ops per LLR update, PM update 2, one compare-exchange of
    the (bitonic-convention) median select 3, one partial-sum XOR per packed 32-bit word.
    (Python twin of pcb_lane_ops_per_frame, kept for the breakdown it returns.)"""
    n = len(mask)
    m = n.bit_length() - 1
    f = g = sa = 0
    ninfo_before = 0
    for i in range(n):
        a = min(L, 1 << min(ninfo_before, 30))
        d = m - 1 if i == 0 else min(m - 1, (i & -i).bit_length() - 1)
        size = 1 << d                      # first recomputed layer has 2^d outputs
        if (i >> d) & 1:
            g += a * size
            f += a * (size - 1)
        else:
            f += a * (2 * size - 1)
        sa += a
        if mask[i] == -1:
            ninfo_before += 1
    ninfo = int((np.asarray(mask) == -1).sum())
    l2 = (2 * L).bit_length() - 1
    ce = ninfo * L * l2 * (l2 + 1) // 2
    x = (f + g) // 2 // 32
    return dict(f=f, g=g, sum_a=sa, ce=ce, xor_words=x, lane_ops=4 * f + 2 * g + 2 * sa + 3 * ce + x)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons DURING the timed region: NVML every ~5 ms, nvidia-smi as a fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self._halt = index, threading.Event()
        self.sm, self.mx, self.reasons, self.power = [], 0.0, set(), 0.0
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        for name, bit in self.REASONS:
            if mask & bit:
                self.reasons.add(name)
        try:
            self.power = max(self.power, n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0)
        except Exception:
            pass

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        r = [c.strip() for c in out.strip().split(",")]
        if len(r) >= 6:
            self.sm.append(float(r[0])); self.mx = max(self.mx, float(r[1]))
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)

    def run(self):
        while not self._halt.is_set():
            try:
                self._sample_nvml() if self.nvml else self._sample_smi()
            except Exception:
                pass
            self._halt.wait(0.005 if self.nvml else 0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "power_w_max": self.power or None,
                "source": "nvml" if self.nvml else "nvidia-smi"}


def cpu_reference_run(frames_total, procs, crc, kind_pref="reference"):
    """Times the reference's own SCL decoder (oracle/_ref) — or the C port when _ref is absent —
    on `procs` host processes (the decoder keeps namespace-global state: processes, not threads).
    Returns (frames/s, kind, seconds, frames)."""
    from oracle.pyoracle import Oracle, Ref
    use_ref = kind_pref == "reference" and Ref.available()
    o = Oracle()
    sigma, inf, mask = o.make_code(N, K, crc, SNR_DB)
    rng = np.random.default_rng(1234)
    per = max(1, frames_total // procs)
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    q = ctx.Queue()

    def work(seed):
        r = np.random.default_rng(seed)
        y = 1.0 + sigma * r.standard_normal((per, N))     # all-zero codeword + AWGN (decode cost is data-typical)
        impl = Ref() if use_ref else Oracle()
        if use_ref:
            impl.lib.ref_scl_init(N, LIST, K, crc)
        t = impl.time_decode(3, N, LIST, K, crc, sigma, y, mask, 1)
        q.put(t)

    ps = [ctx.Process(target=work, args=(int(rng.integers(1 << 30)),)) for _ in range(procs)]
    t0 = time.perf_counter()
    for p in ps:
        p.start()
    times = [q.get() for _ in ps]
    for p in ps:
        p.join()
    wall = time.perf_counter() - t0
    fps = per * procs / max(times)
    return fps, ("reference" if use_ref else "port"), wall, per * procs


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    crc = args.crc
    # bounded sample per step: ~0.75 ms/frame/core
    per_step = max(cores * 64, int(cores * 1300 * 4))
    for _ in range(args.warmup):
        cpu_reference_run(cores * 32, cores, crc)
    t_sum, frames = 0.0, 0
    kind = "port"
    for _ in range(args.steps):
        fps, kind, wall, nf = cpu_reference_run(per_step, cores, crc)
        t_sum += nf / fps
        frames += nf
    fps = frames / t_sum
    gbps = fps * K / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": gbps, "unit": "Gbit/s", "frames_per_s": fps,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_sum / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(crc),
            "sample": {"frames_per_step": per_step},
            "cpu_baseline": {"value": gbps, "unit": "Gbit/s", "cores": cores, "kind": kind,
                             "sample": f"{per_step} distinct frames per step, {cores} processes, decode-only"},
            "e2e": {"value": gbps, "unit": "Gbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---- the other BASELINE.json configurations: device-resident throughput, N = 1 ---------------
def sweep_configs(pb, torch, dev, peaks, traffic):
    cfgs = [  # key, label, decoder, n, k, crc, L, precision, double input, frames
        ("A", "SCDecoderMinsum N=1024 K=512 fp64 (double y)", pb.SCMINSUM, 1024, 512, 0, 1, pb.STRICT, True, 1 << 20),
        ("A_fp32", "SCDecoderMinsum N=1024 K=512 fp32 fast mode (float y; not bit-exact by contract)", pb.SCMINSUM, 1024, 512, 0, 1, pb.FP32, False, 1 << 20),
        ("A_int8", "SCDecoderMinsum N=1024 K=512 8-bit mode, qFraction 2 (double y; a mode of its own: FER in profiles/r2_int8_fer.txt)", pb.SCMINSUM, 1024, 512, 0, 1, pb.int8_q(2), True, 1 << 20),
        ("B", "exact SCDecoder N=256 K=128 fp64 (double y)", pb.SC, 256, 128, 0, 1, pb.STRICT, True, 1 << 20),
        ("C_crc16", "SCLDecoderMinsum L=8 N=1024 K=512 crc=16 (double y)", pb.SCLMINSUM, 1024, 512, 16, 8, pb.STRICT, True, 1 << 20),
        ("conf_txt", "SCLDecoderMinsum L=4 N=2048 K=1024 crc=16 (the reference's shipped conf.txt; double y)", pb.SCLMINSUM, 2048, 1024, 16, 4, pb.STRICT, True, 1 << 18),
        ("D", "SCLDecoderMinsum L=32 N=2048 K=1024 crc=0 (double y)", pb.SCLMINSUM, 2048, 1024, 0, 32, pb.STRICT, True, 1 << 16),
        ("E", "SCDecoderMinsum N=32768 K=16384 fp64 (double y)", pb.SCMINSUM, 32768, 16384, 0, 1, pb.STRICT, True, 1 << 16),
    ]
    out = []
    for key, label, decoder, n, k, crc, L, prec, dbl, B in cfgs:
        torch.cuda.empty_cache()
        code = pb.Code(n, k, crc, SNR_DB)
        dec = pb.Decoder(n, k, crc, L, decoder, prec, device=dev.index)
        dec.set_code(code.mask, code.sigma)
        y = torch.empty((B, n), dtype=torch.float32, device=dev)
        tx = torch.empty((B, dec.words), dtype=torch.int32, device=dev)
        dec.generate_dev(SNR_DB, B, y.data_ptr(), tx.data_ptr(), seed=0x5EED0000, first_frame_id=0)
        if dbl:
            y = y.double()
        sol = torch.empty((B, dec.words), dtype=torch.int32, device=dev)
        for _ in range(2):
            dec.decode_dev(y.data_ptr(), not dbl, sol.data_ptr(), B)
        ms, t_budget = [], time.perf_counter() + 1.5
        while len(ms) < 3 or (time.perf_counter() < t_budget and len(ms) < 50):
            ms.append(dec.decode_dev(y.data_ptr(), not dbl, sol.data_ptr(), B))
        cnt = dec.count_errors_dev(sol.data_ptr(), tx.data_ptr(), B)
        kms = float(np.mean(ms))
        fps = B / (kms * 1e-3)
        ops, upd = pb.lane_ops_per_frame(code.mask, L, decoder)
        peak = dec.alu_peak()
        alg_bytes = n * (8 if dbl else 4) + dec.words * 4
        row = {"config": key, "workload": label, "frames_per_launch": B, "launches_timed": len(ms), "kernel_ms": kms,
               "frames_per_s": fps, "info_gbit_per_s": fps * k / 1e9, "lane_ops_per_frame": ops, "llr_updates_per_frame": upd,
               "alu_peak_tlane_ops": peak / 1e12, "alu_lanes_per_sm": 64 if dec_is_fp64(pb, decoder, prec) else 128,
               "alu_frac": fps * ops / peak, "algorithmic_bytes_per_frame": alg_bytes,
               "hbm_frac_algorithmic": (fps * alg_bytes / 1e9) / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
               "fer": cnt["frame_errors"] / B, "ber": cnt["bit_errors"] / (B * (k if decoder == pb.SCLMINSUM else n))}
        t = (traffic or {}).get(key)
        if t:
            per_frame = (t["dram_bytes_read"] + t["dram_bytes_write"]) / t["frames_per_launch"]
            row["dram_bytes_per_frame"] = per_frame
            row["dram_over_algorithmic"] = per_frame / alg_bytes
            row["traffic_source"] = t.get("source")
        out.append(row)
        dec.close()
        del y, tx, sol
    torch.cuda.empty_cache()
    return out


def encode_leg(pb, torch, dev, peaks):
    """Encoder.cpp:57-83 on packed device-resident frames (pcb_encode_batch_dev): frames/s and the HBM fraction (n / 4 bytes per frame)."""
    out = []
    for key, n, shuffle, B in (("N=1024 natural order (decoder 3)", 1024, 0, 1 << 22), ("N=1024 with ReverseShuffle (decoders 1, 2)", 1024, 1, 1 << 22),
                               ("N=32768 with ReverseShuffle", 32768, 1, 1 << 17)):
        dec = pb.Decoder(n, n // 2, 0, 1, pb.SCMINSUM, device=dev.index)
        u = torch.randint(-2 ** 31, 2 ** 31 - 1, (B, n // 32), dtype=torch.int32, device=dev)
        x = torch.empty_like(u)
        for _ in range(2):
            dec.encode_dev(u.data_ptr(), x.data_ptr(), shuffle, B)
        ms = float(np.mean([dec.encode_dev(u.data_ptr(), x.data_ptr(), shuffle, B) for _ in range(10)]))
        fps = B / (ms * 1e-3)
        gbs = fps * (n / 4) / 1e9
        out.append({"workload": key, "frames_per_launch": B, "kernel_ms": ms, "frames_per_s": fps, "coded_gbit_per_s": fps * n / 1e9,
                    "algorithmic_bytes_per_frame": n // 4, "hbm_GBps": gbs, "hbm_frac": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None})
        dec.close()
        del u, x
    return out


def dec_is_fp64(pb, decoder, prec):
    return decoder == pb.SC or (decoder == pb.SCMINSUM and prec == pb.STRICT)


def montecarlo_leg(pb, dev):
    """pcb_montecarlo: what replaces the reference's loop body (PolarCodes.cpp:132-192) — wall clock of the call."""
    out = []
    for key, decoder, n, k, crc, L, frames in (("C_crc16", pb.SCLMINSUM, 1024, 512, 16, 8, 1 << 24),
                                               ("C_crc0", pb.SCLMINSUM, 1024, 512, 0, 8, 1 << 23),
                                               ("A", pb.SCMINSUM, 1024, 512, 0, 1, 1 << 24)):
        code = pb.Code(n, k, crc, SNR_DB)
        dec = pb.Decoder(n, k, crc, L, decoder, device=dev.index)
        dec.set_code(code.mask, code.sigma)
        dec.montecarlo(SNR_DB, 1 << 20, seed=1)                      # buffers, first launches
        t0 = time.perf_counter()
        r = dec.montecarlo(SNR_DB, frames, seed=0x5EED0000)
        wall = time.perf_counter() - t0
        out.append({"config": key, "frames": r["frames"], "frame_errors": r["frame_errors"], "bit_errors": r["bit_errors"],
                    "fer": r["frame_errors"] / r["frames"], "wall_s": wall, "frames_per_s_wall": r["frames"] / wall,
                    "info_gbit_per_s_wall": r["frames"] / wall * k / 1e9,
                    "frames_per_s_decode_kernels_only": r["frames"] / r["seconds"] if r["seconds"] else None})
        dec.close()
    return out


def main():
    if os.environ.get("PCB_BENCH_WATCHDOG"):                  # debugging aid: where is a hung rank?
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["PCB_BENCH_WATCHDOG"]), exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--crc", type=int, default=0)
    ap.add_argument("--frames", type=int, default=1 << 23,
                    help="frames per step per GPU (SURVEY.md 8(d): >= 2^18 per launch, >= 2 s timed: 2^23 frames ~ 0.2 s)")
    ap.add_argument("--e2e-frames", type=int, default=1 << 20)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the configs sweep and the Monte-Carlo leg")
    ap.add_argument("--kernel-only", action="store_true", help="A/B of kernel variants: the device-resident leg only, short JSON")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.warmup < 3:
        print(f"bench.py: --warmup {args.warmup} < 3 breaks the timing rules (taken as given)", file=sys.stderr)

    import torch
    import torch.distributed as dist
    import polarcodes_b200 as pb

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    code = pb.Code(N, K, args.crc, SNR_DB)
    dec = pb.Decoder(N, K, args.crc, LIST, pb.SCLMINSUM, device=local)
    dec.set_code(code.mask, code.sigma)
    if world > 1:
        # the product's own communicator: rank 0 draws the NCCL id, torch.distributed only carries the 128 bytes
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt = torch.frombuffer(bytearray(pb.comm_unique_id()), dtype=torch.uint8).to(dev)
        dist.broadcast(idt, 0)
        dec.comm_init(bytes(idt.cpu().numpy().tobytes()), rank, world)
    B = args.frames
    words = dec.words

    # synthetic inputs, generated on the device (frame ids disjoint across ranks), widened to the
    # reference's input type (double).  8 KiB/frame * 2^23 = 64 GiB per step: far larger than L2.
    y32 = torch.empty((B, N), dtype=torch.float32, device=dev)
    tx = torch.empty((B, words), dtype=torch.int32, device=dev)
    dec.generate_dev(SNR_DB, B, y32.data_ptr(), tx.data_ptr(), seed=0x5EED0000, first_frame_id=rank * B)
    y64 = torch.empty((B, N), dtype=torch.float64, device=dev)
    for a in range(0, B, 1 << 20):                           # widen in slices: no third full-size temporary
        y64[a:a + (1 << 20)].copy_(y32[a:a + (1 << 20)])
    del y32
    torch.cuda.empty_cache()
    out = torch.empty((B, words), dtype=torch.int32, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        ms = dec.decode_dev(y64.data_ptr(), False, out.data_ptr(), B)         # CUDA events on the launch stream
        return ms, dec.count_errors_dev(out.data_ptr(), tx.data_ptr(), B, allreduce=world > 1)

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    kernel_ms = []
    launches0 = dec.info()["kernel_launches"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ms, counters = step()
        kernel_ms.append(ms)
    barrier()
    elapsed = time.perf_counter() - t0
    gpu_launches = dec.info()["kernel_launches"] - launches0   # prepass + list kernel + (empty) generic follow-up + error count
    clocks = sampler.stop()
    el = torch.tensor([elapsed, float(np.mean(kernel_ms))], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(el, op=dist.ReduceOp.MAX)
    elapsed, kms = float(el[0]), float(el[1])

    if args.kernel_only:
        if rank == 0:
            ops = lane_ops_per_frame(code.mask, LIST)
            info = dec.info()
            peak_lane = info["sm_count"] * 128 * (clocks.get("sm_max_mhz") or 1965.0) * 1e6 / 1e12
            ach = (B / (kms * 1e-3)) * ops["lane_ops"] / 1e12
            print(json.dumps({"frames_per_s": world * B * args.steps / elapsed, "kernel_ms": kms, "roofline": {"frac": ach / peak_lane, "kernel_ms": kms},
                              "clocks": clocks}), flush=True)
        return
    # ---- end to end through the reference-facing call with pinned HOST double buffers -------------------
    Be = min(args.e2e_frames, B)
    hy = torch.empty((Be, N), dtype=torch.float64, pin_memory=True)
    hy.copy_(y64[:Be])
    ref_bits = out[:4096].cpu()
    del y64
    torch.cuda.empty_cache()
    hsol = torch.empty((Be, dec.stride), dtype=torch.int32, pin_memory=True)
    for _ in range(2):
        dec.decode_ptr(hy.data_ptr(), hsol.data_ptr(), Be)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dec.decode_ptr(hy.data_ptr(), hsol.data_ptr(), Be)
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_t[0])
    e2e_fps = world * Be * args.steps / e2e_s
    e2e_ok = bool((torch.from_numpy(pb.unpack_bits(ref_bits.numpy().view(np.uint32), dec.stride)) == hsol[:4096]).all())
    # what the link and the host cores allow: pinned H2D rate of this GPU and the narrowing rate of this rank's
    # thread pool, both measured with every rank busy (pcb_host_probe), combined as the pipeline combines them
    probe = dec.host_probe(hy.data_ptr(), Be)
    barrier()
    # production variant of the same call: float inputs (the decoder works in fp32, SCLDecoderMinsum.cpp:406-408)
    # and packed output bits — 4 KiB in, 64 B out per frame instead of 8 KiB / 2 KiB
    hy32 = torch.empty((Be, N), dtype=torch.float32, pin_memory=True)
    hy32.copy_(hy)
    del hy, hsol
    hbits = torch.empty((Be, words), dtype=torch.int32, pin_memory=True)
    lib = pb.lib()
    for _ in range(2):
        lib.pcb_decode_batch_f32_packed(dec._h, hy32.data_ptr(), hbits.data_ptr(), Be)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        assert lib.pcb_decode_batch_f32_packed(dec._h, hy32.data_ptr(), hbits.data_ptr(), Be) == 0
    barrier()
    e2e32_t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    pr = torch.tensor([probe["h2d_GBps"], probe["narrow_Gvalues_per_s"], probe["bound_frames_per_s"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e32_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(pr, op=dist.ReduceOp.SUM)
    e2e32_fps = world * Be * args.steps / float(e2e32_t[0])
    e2e32_ok = bool((hbits[:4096] == ref_bits).all())
    del hy32, hbits

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        traffic_all = {}
        try:
            traffic_all = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        except Exception:
            pass
        info = dec.info()
        fps = world * B * args.steps / elapsed
        ops = lane_ops_per_frame(code.mask, LIST)
        sm_max = clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
        peak_lane = info["sm_count"] * 128 * sm_max * 1e6 / 1e12         # T lane-ops/s
        ach = (B / (kms * 1e-3)) * ops["lane_ops"] / 1e12
        hbm_bytes = B * (N * 8 + words * 4)
        # DRAM bytes of the dominant kernel per launch, from the committed ncu capture of this workload
        # (scaled to this batch size); algorithmic bytes are hbm_bytes
        traffic, traffic_src, executed = None, None, None
        tj = traffic_all.get("C")
        if tj:
            traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) * B / tj["frames_per_launch"]
            traffic_src = tj.get("source")
            if tj.get("warp_instructions"):
                wi = tj["warp_instructions"] / tj["frames_per_launch"]
                executed = {"warp_instructions_per_frame": wi, "lane_instr_per_frame": 32 * wi,
                            "lane_instr_rate_frac_of_peak": (B / (kms * 1e-3)) * 32 * wi / 1e12 / peak_lane,
                            "note": "executed instructions (ncu smsp__inst_executed.sum) x 32 lanes against the same 128-lane peak; the convention figure above charges 3*CE for a bitonic select the kernel mostly skips"}
        line = {
            "metric": METRIC, "value": fps * K / 1e9, "unit": "Gbit/s", "frames_per_s": fps, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / args.steps,
            "timed_region_s": elapsed,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {**workload_config(args.crc), "frames_per_step_per_gpu": B,
                       "l2": f"inputs larger than L2: {B * N * 8 / 2**30:.0f} GiB of channel values per step and GPU, every frame read once per step; no flush needed"},
            "setup": {"frames_per_step_per_gpu": B,
                      "input": f"double y[B][n] resident in HBM ({B * N * 8 / 2**30:.0f} GiB/step > L2, no flush needed)",
                      "step": "pcb_decode_batch_dev + pcb_count_errors_dev" + (" (ncclAllReduce of the 3 counters inside)" if world > 1 else ""),
                      "collectives_per_step": 1 if world > 1 else 0, "launch": info},
            "roofline": {"bound": "alu", "achieved": ach, "peak": peak_lane, "unit": "Tlane-op/s", "frac": ach / peak_lane,
                         "traffic": traffic, "traffic_source": traffic_src, "lane_ops_per_frame": ops, "kernel_ms": kms,
                         "kernel_ms_covers": "decode_scl8q (narrowing and range check in its prologue) + scl_select + decode_generic (empty redo list), CUDA events on the launch stream",
                         "peak_source": f"{info['sm_count']} SMs x 128 fp32/int lanes x {sm_max:.0f} MHz (SURVEY.md §8(d)); no tensor/HBM bound applies",
                         "executed": executed,
                         "hbm": {"algorithmic_GBps": hbm_bytes / (kms * 1e-3) / 1e9, "peak_GBps": peaks.get("hbm_gbs"),
                                 "frac": (hbm_bytes / (kms * 1e-3) / 1e9) / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None}},
            "e2e": {"value": e2e_fps * K / 1e9, "unit": "Gbit/s", "frames_per_s": e2e_fps,
                    "h2d_bytes_per_step": Be * N * 8, "d2h_bytes_per_step": Be * dec.stride * 4,
                    "frames_per_step_per_gpu": Be, "timed_region_s": e2e_s,
                    "matches_kernel_path": e2e_ok,
                    "input": "pinned host double y[B][n] -> int sol[B][k+crc] (the reference's types)",
                    "bound": {"h2d_GBps_sum": float(pr[0]), "host_narrow_Gvalues_per_s_sum": float(pr[1]),
                              "bound_frames_per_s": float(pr[2]), "frac_of_bound": e2e_fps / float(pr[2]) if float(pr[2]) > 0 else None,
                              "how": "pcb_host_probe on every rank at once: pinned H2D copy rate W and thread-pool double->float rate R; "
                                     "a chunk with head share x narrowed on the host costs max(x/R, (4x + 8(1-x))/W) per value, minimised over x"},
                    "f32_packed_variant": {"value": e2e32_fps * K / 1e9, "unit": "Gbit/s", "frames_per_s": e2e32_fps,
                                           "h2d_bytes_per_step": Be * N * 4, "d2h_bytes_per_step": Be * words * 4,
                                           "matches_kernel_path": e2e32_ok}},
            "gpu_launches": gpu_launches, "clocks": clocks,
            "errors": {"frames": int(counters["frames"]), "frame_errors": int(counters["frame_errors"]),
                       "bit_errors": int(counters["bit_errors"]), "summed_over_ranks_by": "ncclAllReduce (pcb_count_errors_dev)" if world > 1 else "one rank"},
        }
        dec.close()
        if world == 1 and not args.no_extras:
            line["configs"] = sweep_configs(pb, torch, dev, peaks, traffic_all)
            line["montecarlo"] = montecarlo_leg(pb, dev)
            line["encode"] = encode_leg(pb, torch, dev, peaks)
        if not args.no_cpu and world == 1:
            cores = os.cpu_count() or 1
            sample = int(cores * 1300 * 12)
            cfps, kind, wall, nf = cpu_reference_run(sample, cores, args.crc)
            line["cpu_baseline"] = {"value": cfps * K / 1e9, "unit": "Gbit/s", "frames_per_s": cfps, "cores": cores,
                                    "kind": kind, "sample": f"{nf} distinct frames of the same workload, {cores} processes, decode-only, {wall:.1f} s"}
        print(json.dumps(line), flush=True)
    else:
        dec.close()                                           # every rank releases its handle (and its communicator)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
