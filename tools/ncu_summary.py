"""Turn an `ncu --set full` report into the summaries kept under profiles/ (developer tool).

    python tools/ncu_summary.py gpurun_out/r1i/prof.ncu-rep --tag r1i --frames 65536

Writes profiles/<tag>_ncu_full_summary.csv (one column per kernel, the metrics DESIGN.md and
bench.py quote) and profiles/traffic.json (measured DRAM bytes per launch of each stage kernel and
the batch size they were captured at; bench.py scales them to its own batch for
`roofline.traffic`).  Needs the `ncu` CLI, so it runs in the build container on a report fetched
from the GPU box.
"""
from __future__ import annotations

import argparse
import csv
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEEP = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__warps_eligible.avg.per_cycle_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]
# kernel-name fragment -> bench stage; the first match wins (the fused kernel's name contains "cpm_modulate" too)
STAGE = {"cpm_modulate_bank": "modulate_awgn_bank", "cpm_detect": "detect_fused", "encode_frames": "prbs_encode",
         "cpm_modulate": "modulate_awgn", "mf_bank": "mf_bank", "viterbi": "viterbi_bert"}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--tag", required=True)
    ap.add_argument("--frames", type=int, required=True, help="frames per launch in the captured run")
    ap.add_argument("--no-traffic", action="store_true", help="summary only: leave profiles/traffic.json (the bench workload's) alone")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    out = os.path.join(ROOT, "profiles", f"{a.tag}_ncu_full_summary.csv")
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [r[ix["Kernel Name"]].split("(")[0] for r in data])
        for m in KEEP + stall:
            if m in ix:
                w.writerow([m.replace("smsp__average_warps_issue_stalled_", "stall_"), units[ix[m]]] + [r[ix[m]] for r in data])
    traffic = {}
    for r in data:
        name = r[ix["Kernel Name"]]
        for key, stage in STAGE.items():
            if key in name and stage not in traffic:
                tot = 0.0
                for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(r[ix[m]]) * UNIT_SCALE[units[ix[m]]]
                traffic[stage] = {"kernel": name.split("(")[0], "frames": a.frames, "dram_bytes_per_launch": tot,
                                  "source": f"profiles/{a.tag}_ncu_full_summary.csv (ncu --set full, one launch)"}
                if "smsp__inst_executed.sum" in ix:  # warp instructions per launch: the issue-slot roofline's numerator
                    traffic[stage]["warp_instructions_per_launch"] = float(r[ix["smsp__inst_executed.sum"]])
                    traffic[stage]["issue_active_pct"] = float(r[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]])
                break
    if not a.no_traffic:  # merge: stages this capture does not contain keep their earlier entry
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        merged = {}
        if os.path.exists(tpath):
            with open(tpath) as f:
                merged = json.load(f)
        merged.update(traffic)
        with open(tpath, "w") as f:
            json.dump(merged, f, indent=1)
    print(out)
    print(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()
