"""Headless, batched equivalent of the reference's end-to-end script
(examples/soqpsk_detection.py:35-216: PN15 bits -> SOQPSKTrellis4x2DiffEncoded -> cpm_modulate
-> complex AWGN -> PT / PAM matched-filter bank -> SOQPSKTrellisDetector -> error count), run
as the BER waterfall of BASELINE.json configs[3]: Eb/N0 = 0 .. 14 dB, every point simulated
until it has 100 bit errors or the bit budget, frames sharded over the GPUs of the box, the only
collective an NCCL all-reduce of the int64 counter table.

    python examples/soqpsk_ber_waterfall.py                       # one GPU
    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 examples/soqpsk_ber_waterfall.py

Rank 0 prints one JSON object: per point (errors, bits, BER, Clopper-Pearson 95 % interval).
No plots: the reference's figure output is unavailable on this box (no graphics stack).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from waveforms_b200.linksim import LinkConfig, LinkSimulator, ber_confidence_interval  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ebn0", type=float, nargs="+", default=[float(x) for x in range(0, 15)])
    ap.add_argument("--detector", default="PAM", choices=["PAM", "PT"])
    ap.add_argument("--waveform", default="soqpsk-tg")
    ap.add_argument("--n-bits", type=int, default=4096)
    ap.add_argument("--min-errors", type=int, default=100)
    ap.add_argument("--max-bits", type=float, default=1e11)
    ap.add_argument("--batch", type=int, default=65536, help="frames per GPU per kernel launch")
    ap.add_argument("--out", default=None, help="also write the JSON here (rank 0)")
    a = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    sim = LinkSimulator(LinkConfig(waveform=a.waveform, detector=a.detector, n_bits=a.n_bits), device=dev, max_batch=a.batch)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    counts = sim.waterfall(a.ebn0, min_errors=a.min_errors, max_bits=a.max_bits)
    torch.cuda.synchronize(dev)
    seconds = time.perf_counter() - t0

    if rank == 0:
        rows = []
        for eb, (be, se, n) in zip(a.ebn0, counts.tolist()):
            lo, hi = ber_confidence_interval(int(be), int(n))
            rows.append({"ebn0_db": eb, "bit_errors": int(be), "bits": int(n), "ber": be / n if n else None,
                         "ci95": [lo, hi],
                         # which bound retired the point (SURVEY 8d stop rule): enough errors, or the bit budget
                         "stopped_by": "min_errors" if be >= a.min_errors else "max_bits"})
        res = {"waveform": a.waveform, "detector": a.detector, "n_bits": a.n_bits, "n_gpus": world,
               "min_errors": a.min_errors, "max_bits": a.max_bits, "seconds": seconds,
               "simulated_bits": int(np.sum(counts[:, 2])), "bits_per_s": float(np.sum(counts[:, 2]) / seconds),
               "points": rows}
        text = json.dumps(res)
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                f.write(text + "\n")
        sys.stderr.write("Eb/N0  errors         bits        BER        95% interval          stopped by\n")
        for r in rows:
            sys.stderr.write("%5.1f %7d %12.4e %10.3e  [%.3e, %.3e]  %s\n" % (r["ebn0_db"], r["bit_errors"], r["bits"], r["ber"] or 0.0,
                                                                          r["ci95"][0], r["ci95"][1], r["stopped_by"]))
        print(text, flush=True)
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
