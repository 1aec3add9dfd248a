#!/usr/bin/env python
"""bench.py -- Viterbi-decoded SOQPSK-TG bits/s of the link-simulation hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one pass of the whole chain (PRBS+precode -> CPM modulate + AWGN -> PAM
matched-filter bank -> Viterbi + error count) over one batch of synthetic frames per GPU
(BASELINE.json configs[4], largest single-GPU batch: 65536 frames x 4096 PN15 bits,
sps 8, Eb/N0 6 dB).  Weak scaling: every rank runs its own batch of distinct frames;
the only cross-GPU traffic is the final counter all-reduce.  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_BITS, SPS, EBN0_DB, DETECTOR = 4096, 8, 6.0, "PAM"
FRAMES_PER_GPU = 65536
METRIC = "viterbi_decoded_bits_per_s_soqpsk_tg"
# algorithmic bytes per decoded bit (complex64 = 8 B, sps = 8).  The PAM link keeps the
# bank's two raw filter outputs between bank and detector (16 B per call), so the
# bank moves 64 + 16 and the detector 16 + 1/8 bytes per bit (SURVEY's reference-shaped [calls, 3]
# array would be 64 + 24 and 24 + 1/8).
# The fused modulate+AWGN+bank kernel (wf_cpm_modulate_bank, the default link) reads 1 B and writes the
# 16 B of raw bank outputs per bit: the 64 B/bit baseband stays in shared memory.
ALGO_BYTES_PER_BIT = {"prbs_encode": 1.125, "modulate_awgn": 65.0, "mf_bank": 80.0, "viterbi_bert": 16.125,
                      "modulate_awgn_bank": 17.0}
# what ncu shows each stage is limited by (profiles/*_ncu_full_summary.csv)
BOUND = {"prbs_encode": "alu pipe (integer LFSR / precoder work, ~1 B/bit of HBM)",
         "modulate_awgn": "xu pipe + issue (12 MUFU per sample pair: sin/cos of the signal, Box-Muller); HBM write 62 %",
         "mf_bank": "hbm",
         "modulate_awgn_bank": "xu pipe + issue (the modulator's MUFU / Philox work plus the bank's FFMA2 loop from shared "
                               "memory); HBM traffic is 17 B/bit, far from the HBM roof by construction",
         "viterbi_bert": "issue / alu pipe (add-compare-select recursion, one thread per frame)"}


def workload_config(frames, n_gpus):
    return {"workload": f"SOQPSK-TG throughput sweep point: {frames} frames/GPU x {N_BITS} PN15 bits, sps {SPS}, "
                        f"Eb/N0 {EBN0_DB} dB, {DETECTOR} bank, length-2 trellis detector (BASELINE configs[4])",
            "frames_per_gpu": frames, "n_bits": N_BITS, "sps": SPS, "ebn0_db": EBN0_DB, "detector": DETECTOR,
            "parallelism": f"frames sharded over {n_gpus} GPU(s), counter all-reduce only",
            "l2_policy": "per-step working set (z + symbols, ~4.6 GB at 65536 frames; ~24 GB with --unfused) is far larger "
                         "than the 126 MB L2"}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_reference_line(args, as_main):
    """The reference arm / cpu_baseline leg: the oracle port of the reference's numpy chain on
    all host cores, bounded sample of the same workload."""
    from oracle import cpu_bench

    cores = os.cpu_count() or 1
    per_step_s = 6.0 if as_main else 15.0
    if not as_main:
        r = cpu_bench.run_for(per_step_s, n_bits=N_BITS, sps=SPS, ebn0_db=EBN0_DB, detector=DETECTOR, cores=cores)
        return {"value": r["bits_per_s"], "unit": "bits/s", "cores": cores, "kind": "port",
                "sample": f"{r['n_frames']} frames x {N_BITS} bits ({r['seconds']:.1f} s), oracle port of the reference "
                          f"chain, float64, detector vectorised over frames, {cores} processes"}
    pilot = cpu_bench.run(cores * 16, n_bits=N_BITS, sps=SPS, ebn0_db=EBN0_DB, detector=DETECTOR, cores=cores)
    n = max(cores * 16, int(pilot["bits_per_s"] * per_step_s / N_BITS) // (cores * 16) * cores * 16)
    times = []
    for step in range(args.warmup + args.steps):
        r = cpu_bench.run(n, n_bits=N_BITS, sps=SPS, ebn0_db=EBN0_DB, detector=DETECTOR, cores=cores,
                          first_frame=(step + 1) * n)
        if step >= args.warmup:
            times.append(r["seconds"])
    total = sum(times)
    value = n * N_BITS * len(times) / total
    sample = f"{n} frames x {N_BITS} bits per step, oracle port of the reference chain (float64 numpy), {cores} processes"
    cfg = workload_config(FRAMES_PER_GPU, args.gpus)
    cfg["cpu_sample_frames_per_step"] = n
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": "bits/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg, "gpu_launches": 0,
            "cpu_baseline": {"value": value, "unit": "bits/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "bits/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=FRAMES_PER_GPU, help="frames per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fused", action="store_true", help="fused modulate+AWGN+bank kernel (baseband stays in shared memory) instead of the four-kernel link")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries exactly ONE JSON line: anything a library writes to fd 1 on the way
    # (NCCL prints its version banner there) is sent to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(json_fd, (json.dumps(obj) + "\n").encode())

    if args.impl == "reference":
        if rank == 0:
            emit(cpu_reference_line(args, as_main=True))
        return

    # the CPU baseline forks worker processes: take it BEFORE this process owns a CUDA context
    cpu_line = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        cpu_line = cpu_reference_line(args, as_main=False)

    import torch
    import torch.distributed as dist

    from waveforms_b200.linksim import LinkConfig, LinkSimulator
    from waveforms_b200.noise import sigma_from_ebn0_db

    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize(dev)

    B, K, W = args.frames, args.steps, max(args.warmup, 3)
    sim = LinkSimulator(LinkConfig(n_bits=N_BITS, sps=SPS, detector=DETECTOR, extra={"fused": args.fused}),
                        device=dev, max_batch=B)
    KERNELS = list(sim.stage_names)
    NK = len(KERNELS)
    sigma = torch.full((B,), float(sigma_from_ebn0_db(EBN0_DB, SPS)), dtype=torch.float32, device=dev)
    point = torch.zeros((B,), dtype=torch.int32, device=dev)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=dev)
    frame_of = lambda step: (step * world + rank) * B  # noqa: E731  distinct frames per rank and step

    # ---------------- device-resident loop: `value` ------------------------------------
    for s in range(W):
        sim.run_batch(frame_of(s), sigma, point, counters)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    launches0 = sim.launches
    events = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    counters.zero_()
    barrier()
    t0.record(torch.cuda.current_stream(dev))
    for s in range(K):
        sim.run_batch(frame_of(W + s), sigma, point, counters, events=events)
    t1.record(torch.cuda.current_stream(dev))
    barrier()
    clocks = sampler.stop() if sampler else None
    elapsed_ms = t0.elapsed_time(t1)
    launches = sim.launches - launches0
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(counters, op=dist.ReduceOp.SUM)  # the path's one collective
    elapsed_ms = float(t.item())
    bit_err, _, compared = counters[0].tolist()
    per_kernel_ms = [statistics.fmean(events[(NK + 1) * s + k].elapsed_time(events[(NK + 1) * s + k + 1]) for s in range(K))
                     for k in range(NK)]

    # ---------------- end-to-end loop through the public host call: `e2e` ---------------------
    for s in range(2):
        sim.run([EBN0_DB], B, first_frame=frame_of(s))
    barrier()
    w0 = time.perf_counter()
    for s in range(K):
        host_counts = sim.run([EBN0_DB], B, first_frame=frame_of(W + s))  # pinned H2D in, D2H counters out
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - w0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())

    if rank == 0:
        bits_per_step = world * B * N_BITS
        value = bits_per_step * K / (elapsed_ms * 1e-3)
        peak, peak_src = peaks()
        dom = max(range(NK), key=lambda k: per_kernel_ms[k])
        name = KERNELS[dom]
        algo_bytes = ALGO_BYTES_PER_BIT[name] * B * N_BITS
        achieved = algo_bytes / (per_kernel_ms[dom] * 1e-3) / 1e9
        # measured DRAM bytes per launch (ncu --set full, tools/ncu_summary.py), scaled to this batch
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        tj_all = {}
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj_all = json.load(f)

        def traffic_of(stage):
            tj = tj_all.get(stage)
            return tj["dram_bytes_per_launch"] * (B / tj["frames"]) if tj and tj.get("frames") else None

        traffic = traffic_of(name)
        line = {
            "metric": METRIC, "value": value, "unit": "bits/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": elapsed_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(B, world),
            "gpu_launches": launches,
            "e2e": {"value": bits_per_step * K / e2e_s, "unit": "bits/s", "h2d_bytes_per_step": B * 8,
                    "d2h_bytes_per_step": int(host_counts.nbytes)},
            "roofline": {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_bit": ALGO_BYTES_PER_BIT[name], "binding_resource": BOUND[name]},
            "link": "fused (prbs | modulate+awgn+bank | viterbi+bert)" if sim.fused else "four kernels",
            "kernels": {KERNELS[k]: {"ms": per_kernel_ms[k],
                                     "algo_gbs": ALGO_BYTES_PER_BIT[KERNELS[k]] * B * N_BITS / (per_kernel_ms[k] * 1e-3) / 1e9,
                                     "hbm_frac": ALGO_BYTES_PER_BIT[KERNELS[k]] * B * N_BITS / (per_kernel_ms[k] * 1e-3) / 1e9 / peak,
                                     "algorithmic_bytes": ALGO_BYTES_PER_BIT[KERNELS[k]] * B * N_BITS,
                                     "traffic": traffic_of(KERNELS[k]), "bound": BOUND[KERNELS[k]],
                                     "share": per_kernel_ms[k] / sum(per_kernel_ms)} for k in range(NK)},
            "clocks": clocks,
            "ber": {"ebn0_db": EBN0_DB, "bit_errors": bit_err, "compared": compared,
                    "ber": bit_err / compared if compared else None},
        }
        if cpu_line is not None:
            line["cpu_baseline"] = cpu_line
        emit(line)
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
