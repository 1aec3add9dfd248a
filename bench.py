#!/usr/bin/env python
"""bench.py -- Viterbi-decoded SOQPSK-TG bits/s of the link-simulation hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--unfused]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --sweep 1024,4096,...,1048576 [--sweep-out profiles/r2_sweep_n1.json]   (BASELINE configs[4])

One step = one pass of the whole chain (PRBS+precode -> CPM modulate + AWGN -> PAM
matched-filter bank -> Viterbi + error count) over one batch of synthetic frames per GPU
(BASELINE.json configs[4], largest single-GPU batch: 65536 frames x 4096 PN15 bits,
sps 8, Eb/N0 6 dB).  Weak scaling: every rank runs its own batch of distinct frames;
the only cross-GPU traffic is the final counter all-reduce.  Rank 0 prints ONE JSON line.

--sweep runs BASELINE configs[4] as written: a FIXED TOTAL of B frames, B in
{2^10 .. 2^20}, split over the N ranks (strong scaling), each point timed (a) device-resident and
(b) through LinkSimulator.sweep() with host inputs / outputs and the NCCL counter all-reduce inside
the timed region, so that the launch-latency floor of small batches is visible, not hidden.
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
SM_COUNT = 148  # B200
FRAMES_PER_GPU = 65536
METRIC = "viterbi_decoded_bits_per_s_soqpsk_tg"
# algorithmic bytes per decoded bit (complex64 = 8 B, sps = 8).  The PAM link keeps the
# bank's two raw filter outputs between bank and detector (16 B per call), so the
# bank moves 64 + 16 and the detector 16 + 1/8 bytes per bit (SURVEY's reference-shaped [calls, 3]
# array would be 64 + 24 and 24 + 1/8).
# The fused modulate+AWGN+bank kernel (wf_cpm_modulate_bank, chosen with --fused / --unfused, see DEFAULT_FUSED)
# reads 1 B and writes the 16 B of raw bank outputs per bit: the 64 B/bit baseband stays in shared memory.
ALGO_BYTES_PER_BIT = {"prbs_encode": 1.125, "modulate_awgn": 65.0, "mf_bank": 80.0, "viterbi_bert": 16.125,
                      "modulate_awgn_bank": 17.0}
# what ncu shows each stage is limited by (profiles/*_ncu_full_summary.csv)
DEFAULT_FUSED = True  # the fused link is the faster one since round 2 (warp-autonomous kernel) and bit-identical
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


def unmodified_reference_line():
    """VERDICT r1 #9 / BASELINE.md section 3: the UNMODIFIED reference (baseline/_ref, staged by
    __graft_entry__.build()) timed on this box's host cores through oracle/ref_bench.py -- single process (upstream is
    single-threaded) and multiprocessing.Pool(all cores) over independent frames; a few frames each (~0.3 s per
    frame per core), so ~10 s in all.  Reported beside cpu_baseline (the vectorised oracle port, which is ~4x faster on
    the same cores and therefore the conservative denominator); says "unavailable" when the package is not staged."""
    try:
        from oracle import ref_bench
        if ref_bench.locate() is None:
            return {"unavailable": "baseline/_ref not staged (run __graft_entry__.build() in the build container)"}
        r = ref_bench.measure(frames_single=4, frames_pool_per_core=4, ebn0=EBN0_DB, detector=DETECTOR)
        return {"kind": "reference", "unit": "bits/s", "cores": r["cores"], "cpu_model": r["cpu_model"],
                "single_process": r["single_process"]["bits_per_s"], "pool_all_cores": r["pool_all_cores"]["bits_per_s"],
                "sample": f"{r['single_process']['frames']} frames single-process, {r['pool_all_cores']['frames']} frames on "
                          f"{r['cores']} processes, {N_BITS} bits each, Eb/N0 {EBN0_DB} dB, {DETECTOR} bank; the reference's own "
                          "functions driven by the headless replay of examples/soqpsk_detection.py:45-209"}
    except Exception as e:  # noqa: BLE001  (a baseline leg must never take the bench line down)
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}


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


def sweep_mode(args, sim_factory, dev, rank, world, local_rank, barrier, emit):
    """BASELINE configs[4]: fixed TOTAL frame counts split over the ranks (strong scaling)."""
    import torch
    import torch.distributed as dist

    from waveforms_b200.linksim import shard_range
    from waveforms_b200.noise import sigma_from_ebn0_db

    totals = [int(x) for x in args.sweep.split(",")]
    rows = []
    sigma_v = float(sigma_from_ebn0_db(EBN0_DB, SPS))
    sim = None
    for total in totals:
        lo, hi = shard_range(total, rank, world)
        n_loc = hi - lo
        mb = max(1, min(n_loc, FRAMES_PER_GPU))
        if sim is None or sim.max_batch != mb:
            sim = None
            torch.cuda.empty_cache()
            sim = sim_factory(mb)
        sigma = torch.full((mb,), sigma_v, dtype=torch.float32, device=dev)
        point = torch.zeros((mb,), dtype=torch.int32, device=dev)
        counters = torch.zeros((1, 3), dtype=torch.int64, device=dev)

        def one_pass(base):
            for f0 in range(0, n_loc, mb):
                b = min(mb, n_loc - f0)
                sim.run_batch(base + lo + f0, sigma[:b], point[:b], counters)

        reps = max(3, min(200, int(3e5 // max(total, 1))))
        for w in range(3):
            one_pass(w * total)
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(torch.cuda.current_stream(dev))
        for r in range(reps):
            one_pass((3 + r) * total)
        t1.record(torch.cuda.current_stream(dev))
        barrier()
        t = torch.tensor([t0.elapsed_time(t1) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
        # end to end: host Eb/N0 in, host counters out, counter all-reduce inside
        for w in range(2):
            sim.sweep([EBN0_DB], total, first_frame=w * total)
        barrier()
        w0 = time.perf_counter()
        for r in range(reps):
            sim.sweep([EBN0_DB], total, first_frame=(2 + r) * total)
        torch.cuda.synchronize(dev)
        te = torch.tensor([(time.perf_counter() - w0) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_ms = float(te.item()) * 1e3
        rows.append({"total_frames": total, "n_gpus": world, "frames_per_gpu": n_loc if world == 1 else [total // world, -(-total // world)],
                     "reps": reps, "device_ms": dev_ms, "device_bits_per_s": total * N_BITS / (dev_ms * 1e-3),
                     "e2e_ms": e2e_ms, "e2e_bits_per_s": total * N_BITS / (e2e_ms * 1e-3),
                     "launches_per_pass": len(sim.stage_names) * -(-max(n_loc, 1) // mb)})
    if rank == 0:
        out = {"metric": METRIC, "unit": "bits/s", "n_gpus": world, "scaling": "strong", "link": "fused" if sim.fused else "four kernels",
               "config": {"workload": f"BASELINE configs[4] throughput sweep: total frames x {N_BITS} PN15 bits, sps {SPS}, "
                                      f"Eb/N0 {EBN0_DB} dB, {DETECTOR} bank; total frames split over {world} GPU(s)"},
               "points": rows}
        if args.sweep_cpu and world == 1:
            from oracle import cpu_bench
            cores = os.cpu_count() or 1
            cpu = []
            for total in totals:
                if total <= args.sweep_cpu:
                    r = cpu_bench.run(total, n_bits=N_BITS, sps=SPS, ebn0_db=EBN0_DB, detector=DETECTOR, cores=cores)
                    cpu.append({"total_frames": total, "seconds": r["seconds"], "bits_per_s": r["bits_per_s"], "measured": True})
                elif cpu:
                    cpu.append({"total_frames": total, "seconds": total * N_BITS / cpu[-1]["bits_per_s"],
                                "bits_per_s": cpu[-1]["bits_per_s"], "measured": False})
            out["cpu_port"] = {"cores": cores, "kind": "port", "points": cpu,
                               "note": "oracle port on all host cores; points beyond --sweep-cpu frames are the last measured rate extrapolated"}
        if args.sweep_out:
            os.makedirs(os.path.dirname(os.path.abspath(args.sweep_out)), exist_ok=True)
            with open(args.sweep_out, "w") as f:
                json.dump(out, f, indent=1)
        emit(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=FRAMES_PER_GPU, help="frames per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fused", action="store_true", help="fused modulate+AWGN+bank kernel (baseband stays in shared memory)")
    ap.add_argument("--unfused", action="store_true", help="four-kernel link (baseband through HBM)")
    ap.add_argument("--sweep", default=None, help="comma-separated TOTAL frame counts: BASELINE configs[4], strong scaling")
    ap.add_argument("--sweep-out", default=None, help="also write the sweep result to this JSON file")
    ap.add_argument("--sweep-cpu", type=int, default=0, help="time the CPU port for totals up to this many frames (N=1 only)")
    args = ap.parse_args()
    fused = (DEFAULT_FUSED or args.fused) and not args.unfused
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
    cpu_line = ref_line = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline and not args.sweep:
        cpu_line = cpu_reference_line(args, as_main=False)
        ref_line = unmodified_reference_line()

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

    def sim_factory(max_batch):
        return LinkSimulator(LinkConfig(n_bits=N_BITS, sps=SPS, detector=DETECTOR, extra={"fused": fused}),
                             device=dev, max_batch=max_batch)

    if args.sweep:
        sweep_mode(args, sim_factory, dev, rank, world, local_rank, barrier, emit)
        if world > 1:
            dist.barrier(device_ids=[local_rank])
            dist.destroy_process_group()
        return

    B, K, W = args.frames, args.steps, max(args.warmup, 3)
    sim = sim_factory(B)
    KERNELS = list(sim.stage_names)
    NK = len(KERNELS)
    sigma = torch.full((B,), float(sigma_from_ebn0_db(EBN0_DB, SPS)), dtype=torch.float32, device=dev)
    point = torch.zeros((B,), dtype=torch.int32, device=dev)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=dev)
    frame_of = lambda step: (step * world + rank) * B  # noqa: E731  distinct frames per rank and step

    # ---------------- device-resident loop: `value` (no events inside the timed region) ----------
    for s in range(W):
        sim.run_batch(frame_of(s), sigma, point, counters)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    launches0 = sim.launches
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    counters.zero_()
    barrier()
    t0.record(torch.cuda.current_stream(dev))
    for s in range(K):
        sim.run_batch(frame_of(W + s), sigma, point, counters)
    t1.record(torch.cuda.current_stream(dev))
    barrier()
    elapsed_ms = t0.elapsed_time(t1)
    launches = sim.launches - launches0
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(counters, op=dist.ReduceOp.SUM)  # the path's one collective
    elapsed_ms = float(t.item())
    bit_err, _, compared = counters[0].tolist()

    # ---------------- per-kernel pass: the same K steps again with CUDA events between the launches
    # (same stream, clocks still sampled) -- kept OUT of the region `value` is timed over
    events = []
    scratch = torch.zeros_like(counters)
    for s in range(K):
        sim.run_batch(frame_of(W + s), sigma, point, scratch, events=events)
    barrier()
    clocks = sampler.stop() if sampler else None
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
        # measured DRAM bytes per launch: NOT measured by this run (ncu cannot run inside the timed bench);
        # taken from profiles/traffic.json = one `ncu --set full` launch per stage kernel captured on the same
        # tree by tools/ncu_summary.py at `frames` frames, scaled linearly to this batch (traffic_source says so)
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        tj_all = {}
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj_all = json.load(f)

        def traffic_of(stage):
            tj = tj_all.get(stage)
            return tj["dram_bytes_per_launch"] * (B / tj["frames"]) if tj and tj.get("frames") else None

        def traffic_src(stage):
            tj = tj_all.get(stage)
            if not (tj and tj.get("frames")):
                return None
            return (f"profiles/traffic.json: ncu --set full, one launch at {tj['frames']} frames "
                    f"({tj.get('source', 'source?')}), scaled x{B / tj['frames']:g} to this batch")

        def issue_roof(stage, ms, sm_mhz):
            """Issue-slot roofline of an instruction-bound kernel: warp instructions per launch (ncu
            smsp__inst_executed.sum of the same capture traffic.json comes from, scaled to this batch) over the
            kernel's live duration, against 148 SMs x 4 schedulers x 1 warp instruction per clock at the SM clock
            sampled during this run."""
            tj = tj_all.get(stage)
            if not (tj and tj.get("warp_instructions_per_launch") and sm_mhz):
                return None
            inst = tj["warp_instructions_per_launch"] * (B / tj["frames"])
            peak_ips = SM_COUNT * 4 * sm_mhz * 1e6
            return {"warp_instructions": inst, "achieved_per_s": inst / (ms * 1e-3), "peak_per_s": peak_ips,
                    "frac": inst / (ms * 1e-3) / peak_ips, "ncu_issue_active_pct": tj.get("issue_active_pct"),
                    "source": tj.get("source")}

        traffic = traffic_of(name)
        line = {
            "metric": METRIC, "value": value, "unit": "bits/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": elapsed_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(B, world),
            "gpu_launches": launches,
            "e2e": {"value": bits_per_step * K / e2e_s, "unit": "bits/s", "h2d_bytes_per_step": B * 8,
                    "d2h_bytes_per_step": int(host_counts.nbytes)},
            "roofline": {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src(name),
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_bit": ALGO_BYTES_PER_BIT[name], "binding_resource": BOUND[name],
                         "issue": issue_roof(name, per_kernel_ms[dom], clocks.get("sm_mhz") if clocks else None)},
            "link": "fused (prbs | modulate+awgn+bank | viterbi+bert)" if sim.fused else "four kernels",
            "kernels": {KERNELS[k]: {"ms": per_kernel_ms[k],
                                     "algo_gbs": ALGO_BYTES_PER_BIT[KERNELS[k]] * B * N_BITS / (per_kernel_ms[k] * 1e-3) / 1e9,
                                     "hbm_frac": ALGO_BYTES_PER_BIT[KERNELS[k]] * B * N_BITS / (per_kernel_ms[k] * 1e-3) / 1e9 / peak,
                                     "algorithmic_bytes": ALGO_BYTES_PER_BIT[KERNELS[k]] * B * N_BITS,
                                     "traffic": traffic_of(KERNELS[k]), "bound": BOUND[KERNELS[k]],
                                     "issue": issue_roof(KERNELS[k], per_kernel_ms[k], clocks.get("sm_mhz") if clocks else None),
                                     "share": per_kernel_ms[k] / sum(per_kernel_ms)} for k in range(NK)},
            "per_kernel_timing": "separate pass of the same K steps with CUDA events between launches (not inside the region `value` is timed over)",
            "clocks": clocks,
            "ber": {"ebn0_db": EBN0_DB, "bit_errors": bit_err, "compared": compared,
                    "ber": bit_err / compared if compared else None},
        }
        if cpu_line is not None:
            line["cpu_baseline"] = cpu_line
        if ref_line is not None:
            line["cpu_baseline_reference"] = ref_line
        emit(line)
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
