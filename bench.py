#!/usr/bin/env python
"""bench.py — headline benchmark of the polar decode hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the decoder over one batch of synthetic frames (on-device Philox
BPSK/AWGN codewords of the AWGN-constructed code, 2.0 dB).  Workload: SCLDecoderMinsum L=8,
N=1024, K=512 (BASELINE.json configs[2], the configuration the metric is quoted on).
  value     decoded info Gbit/s with the channel values already resident in HBM (kernel path)
  e2e       same metric through the reference-facing C-ABI call pcb_decode_batch with HOST
            (pinned) double buffers: H2D + kernel + D2H inside the timed region
  roofline  ALU lane-ops (SURVEY.md §8(d) convention, counted exactly from the frozen mask)
            / kernel time from CUDA events / (SMs * 128 lanes * max SM clock)
  cpu_baseline  the unmodified reference (oracle/_ref) or the C port timed on the host cores
Multi-GPU: one process per GPU (torchrun), frames sharded, no data-path collective; NCCL
only sums the error counters (weak scaling).
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


def lane_ops_per_frame(mask, L):
    """f=4, g=2 lane-ops per LLR update, PM update 2, one compare-exchange of
    the (bitonic-convention) median select 3, one partial-sum XOR per packed 32-bit word."""
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
    """Samples SM clock and throttle reasons DURING the timed region: NVML every ~2 ms (the timed
    region of a default run is only ~50 ms), nvidia-smi as a fallback."""
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
            self._halt.wait(0.002 if self.nvml else 0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "power_w_max": self.power or None,
                "source": "nvml" if self.nvml else "nvidia-smi"}


def cpu_reference_run(frames_total, procs, crc, kind_pref="reference"):
    """Times the reference's own SCL decoder (oracle/_ref) — or the C port when _ref is absent —
    on `procs` host processes (the decoder keeps namespace-global state: processes, not threads).
    Returns (frames/s, kind, seconds)."""
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
            "config": {"workload": f"SCLDecoderMinsum L={LIST} N={N} K={K} crc={crc} AWGN-construction {SNR_DB} dB",
                       "frames_per_step": per_step},
            "cpu_baseline": {"value": gbps, "unit": "Gbit/s", "cores": cores, "kind": kind,
                             "sample": f"{per_step} distinct frames per step, {cores} processes, decode-only"},
            "e2e": {"value": gbps, "unit": "Gbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--crc", type=int, default=0)
    ap.add_argument("--frames", type=int, default=1 << 18, help="frames per step per GPU (SURVEY.md 8(d): >= 2^18 per launch)")
    ap.add_argument("--e2e-frames", type=int, default=1 << 16)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import polarcodes_b200 as pb

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    code = pb.Code(N, K, args.crc, SNR_DB)
    dec = pb.Decoder(N, K, args.crc, LIST, pb.SCLMINSUM, device=local)
    dec.set_code(code.mask, code.sigma)
    B = args.frames
    words = dec.words

    # synthetic inputs, generated on the device (frame ids disjoint across ranks), widened to the
    # reference's input type (double).  8 KiB/frame * 131072 = 1 GiB per step: far larger than L2.
    y32 = torch.empty((B, N), dtype=torch.float32, device=dev)
    tx = torch.empty((B, words), dtype=torch.int32, device=dev)
    dec.generate_dev(SNR_DB, B, y32.data_ptr(), tx.data_ptr(), seed=0x5EED0000, first_frame_id=rank * B)
    y64 = y32.double()
    del y32
    out = torch.empty((B, words), dtype=torch.int32, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        dec.decode_dev(y64.data_ptr(), False, out.data_ptr(), B)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    kernel_ms = []
    launches0 = dec.info()["kernel_launches"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        kernel_ms.append(dec.decode_dev(y64.data_ptr(), False, out.data_ptr(), B))   # CUDA events on the launch stream
    barrier()
    elapsed = time.perf_counter() - t0
    gpu_launches = dec.info()["kernel_launches"] - launches0   # tuned kernel + its (normally empty) generic follow-up
    clocks = sampler.stop()
    # error counters of the last step, summed over ranks with NCCL (the path's only collective)
    kmask = torch.ones(words * 32, dtype=torch.bool, device=dev)
    diff = (out ^ tx)
    bits = ((diff.unsqueeze(-1) >> torch.arange(32, device=dev, dtype=torch.int32)) & 1).reshape(B, -1)[:, :K]
    counters = torch.stack([torch.tensor(B, device=dev, dtype=torch.int64), bits.any(dim=1).sum(), bits.sum()])
    el = torch.tensor([elapsed, float(np.mean(kernel_ms))], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(counters, op=dist.ReduceOp.SUM)
        dist.all_reduce(el, op=dist.ReduceOp.MAX)
    elapsed, kms = float(el[0]), float(el[1])
    del kmask, diff, bits

    # end-to-end through the reference-facing call with pinned HOST double buffers
    Be = min(args.e2e_frames, B)
    hy = torch.empty((Be, N), dtype=torch.float64).pin_memory()
    hy.copy_(y64[:Be])
    hsol = torch.empty((Be, dec.stride), dtype=torch.int32).pin_memory()
    del y64
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
    e2e_fps = world * Be * args.steps / float(e2e_t[0])
    # e2e output equals the kernel-path output (same frames)
    e2e_ok = bool((torch.from_numpy(pb.unpack_bits(out[:256].cpu().numpy().view(np.uint32), dec.stride)) == hsol[:256]).all())
    # production variant of the same call: float inputs (the decoder works in fp32, SCLDecoderMinsum.cpp:406-408)
    # and packed output bits — 4 KiB in, 64 B out per frame instead of 8 KiB / 2 KiB
    hy32 = torch.empty((Be, N), dtype=torch.float32).pin_memory()
    hy32.copy_(hy)
    hbits = torch.empty((Be, words), dtype=torch.int32).pin_memory()
    lib = pb.lib()
    for _ in range(2):
        lib.pcb_decode_batch_f32_packed(dec._h, hy32.data_ptr(), hbits.data_ptr(), Be)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        assert lib.pcb_decode_batch_f32_packed(dec._h, hy32.data_ptr(), hbits.data_ptr(), Be) == 0
    barrier()
    e2e32_t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e32_t, op=dist.ReduceOp.MAX)
    e2e32_fps = world * Be * args.steps / float(e2e32_t[0])
    e2e32_ok = bool((hbits[:256] == out[:256].cpu()).all())

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        info = dec.info()
        fps = world * B * args.steps / elapsed
        ops = lane_ops_per_frame(code.mask, LIST)
        sm_max = clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
        peak_lane = info["sm_count"] * 128 * sm_max * 1e6 / 1e12         # T lane-ops/s
        ach = (B / (kms * 1e-3)) * ops["lane_ops"] / 1e12
        hbm_bytes = B * (N * 8 + words * 4)
        # DRAM bytes of the dominant kernel per launch, from the committed ncu capture of this
        # command (scaled to this batch size); algorithmic bytes are hbm_bytes
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r1_scl8q_traffic.json")))
            traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) * B / tj["frames_per_launch"]
            traffic_src = tj["source"]
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": fps * K / 1e9, "unit": "Gbit/s", "frames_per_s": fps, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * elapsed / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"SCLDecoderMinsum L={LIST} N={N} K={K} crc={args.crc} AWGN-construction {SNR_DB} dB",
                       "frames_per_step_per_gpu": B, "input": f"double y[B][n] resident in HBM ({B * N * 8 / 2**30:.0f} GiB/step > L2, no flush needed)",
                       "launch": info},
            "roofline": {"bound": "alu", "achieved": ach, "peak": peak_lane, "unit": "Tlane-op/s", "frac": ach / peak_lane,
                         "traffic": traffic, "traffic_source": traffic_src, "lane_ops_per_frame": ops, "kernel_ms": kms,
                         "kernel_ms_covers": "scl8_prepass + decode_scl8q + decode_generic (empty redo list), CUDA events on the launch stream",
                         "peak_source": f"{info['sm_count']} SMs x 128 fp32/int lanes x {sm_max:.0f} MHz (SURVEY.md §8(d)); no tensor/HBM bound applies",
                         "hbm": {"algorithmic_GBps": hbm_bytes / (kms * 1e-3) / 1e9, "peak_GBps": peaks.get("hbm_gbs"),
                                 "frac": (hbm_bytes / (kms * 1e-3) / 1e9) / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None}},
            "e2e": {"value": e2e_fps * K / 1e9, "unit": "Gbit/s", "frames_per_s": e2e_fps,
                    "h2d_bytes_per_step": Be * N * 8, "d2h_bytes_per_step": Be * dec.stride * 4,
                    "matches_kernel_path": e2e_ok,
                    "input": "pinned host double y[B][n] -> int sol[B][k+crc] (the reference's types)",
                    "f32_packed_variant": {"value": e2e32_fps * K / 1e9, "unit": "Gbit/s", "frames_per_s": e2e32_fps,
                                           "h2d_bytes_per_step": Be * N * 4, "d2h_bytes_per_step": Be * words * 4,
                                           "matches_kernel_path": e2e32_ok}},
            "gpu_launches": gpu_launches, "clocks": clocks,
            "errors": {"frames": int(counters[0]), "frame_errors": int(counters[1]), "bit_errors": int(counters[2])},
        }
        if not args.no_cpu:
            cores = os.cpu_count() or 1
            sample = int(cores * 1300 * 12)
            cfps, kind, wall, nf = cpu_reference_run(sample, cores, args.crc)
            line["cpu_baseline"] = {"value": cfps * K / 1e9, "unit": "Gbit/s", "frames_per_s": cfps, "cores": cores,
                                    "kind": kind, "sample": f"{nf} distinct frames of the same workload, {cores} processes, decode-only, {wall:.1f} s"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
