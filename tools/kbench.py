"""Per-kernel timing of the SOQPSK-TG chain at a chosen batch (developer tool, GPU box only).

    python tools/kbench.py [--frames 16384] [--reps 5] [--env WF_MOD_TUNE=0,1,2]

Times each of the four stage kernels with CUDA events around `reps` back-to-back batches and
prints ms per batch + algorithmic GB/s.  `--env NAME=v0,v1,...` repeats the measurement with an
environment knob set to each value (used to compare kernel variants in one gpurun call).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from waveforms_b200.linksim import LinkConfig, LinkSimulator  # noqa: E402

BYTES = {"prbs_encode": 1.125, "modulate_awgn": 65.0, "mf_bank": 80.0, "viterbi_bert": 16.125,  # PAM raw-bank link (bench.py)
         "modulate_awgn_bank": 17.0, "detect_fused": 65.0}  # detect_fused: 64 B of rx + 1 B of transmitted symbol per symbol


def measure(sim, B, reps, cfg):
    dev = sim.device
    sigma = (cfg.sps / (2.0 * sim.bits_per_symbol * 10.0 ** 0.6)) ** 0.5  # 6 dB
    sg = torch.full((B,), sigma, dtype=torch.float32, device=dev)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=dev)
    for _ in range(2):
        sim.run_batch(0, sg, None, counters)
    torch.cuda.synchronize(dev)
    tot = dict.fromkeys(sim.stage_names, 0.0)
    for r in range(reps):
        ev = []
        sim.run_batch(r * B, sg, None, counters, events=ev)
        torch.cuda.synchronize(dev)
        for k, name in enumerate(sim.stage_names):
            tot[name] += ev[k].elapsed_time(ev[k + 1])
    bits = B * cfg.n_bits
    out = {}
    for name, ms in tot.items():
        ms /= reps
        out[name] = {"ms": round(ms, 4), "algo_gbs": round(BYTES[name] * bits / ms / 1e6, 1)}
    out["waveform"] = cfg.waveform
    out["extra"] = {k: v for k, v in cfg.extra.items() if v}
    out["frames"] = B
    out["total_ms"] = round(sum(out[name]["ms"] for name in sim.stage_names), 4)
    out["bits_per_s"] = round(bits / out["total_ms"] * 1e3, 1)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=16384)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--detector", default="PAM")
    ap.add_argument("--waveform", default="soqpsk-tg", help="LinkConfig.waveform (algo_gbs figures are SOQPSK/PAM bytes)")
    ap.add_argument("--env", default=None)
    ap.add_argument("--fused", action="store_true", help="(the fused link is the default; kept for old scripts)")
    ap.add_argument("--unfused", action="store_true", help="four-kernel SOQPSK link / unfused CPM detector")
    ap.add_argument("--extra", default=None, help="JSON dict merged into LinkConfig.extra (e.g. '{\"tune\": 4}')")
    a = ap.parse_args()
    extra = {"fused": True} if a.fused else {}
    if a.unfused:
        extra.update({"fused": False, "unfused_detector": True})
    if a.extra:
        extra.update(json.loads(a.extra))
    cfg = LinkConfig(n_bits=4096, detector=a.detector, waveform=a.waveform, extra=extra)
    sim = LinkSimulator(cfg, device=torch.device("cuda", 0), max_batch=a.frames)
    if a.env:
        name, vals = a.env.split("=")
        for v in vals.split(","):
            os.environ[name] = v
            print(json.dumps({"env": f"{name}={v}", **measure(sim, a.frames, a.reps, cfg)}), flush=True)
    else:
        print(json.dumps(measure(sim, a.frames, a.reps, cfg)), flush=True)


if __name__ == "__main__":
    main()
