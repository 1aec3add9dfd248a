"""Regenerate the reference BER waterfall of BASELINE.md section 2c by running the
REFERENCE ITSELF (read-only at /root/reference): SOQPSK-TG, sps 8, N=4096, Eb/N0 0..12 dB,
--frames per point (committed table: 1024), frame f = PN15 bits [f*4096, ...) mod 32767, noise
PCG64(1000*EbN0_dB + f), both detectors on the same received frame.

    python oracle/make_ber_table.py [--frames 1024] [--procs 8]  ->  tests/golden/g10_ber_table.npz

Run once in the build container; the output is committed.  Test infrastructure only."""
from __future__ import annotations

import argparse
import multiprocessing as mp
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


def _one(args):
    ebn0, f = args
    import make_golden as mg  # imports the reference from /root/reference

    sps = 8
    seq = np.array(mg.PNSequence(15).generate_sequence(), dtype=np.uint8)
    bits = seq[(f * 4096 + np.arange(4096)) % seq.size]
    sigma = float(np.sqrt(sps / (2 * 10 ** (ebn0 / 10))))
    rng = np.random.Generator(np.random.PCG64(seed=1000 * ebn0 + f))
    r = mg.replay_frame(bits, mg.freq_pulse_soqpsk_tg(sps), sps, sigma, rng, {"PT": 0, "PAM": 0, "PT@-1": -1})
    return (ebn0, r["PT"]["bit_errors"], r["PAM"]["bit_errors"], r["PAM"]["compared"],
            r["PT@-1"]["bit_errors"], r["PT@-1"]["compared"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--out", default=os.path.join(HERE, "..", "tests", "golden", "g10_ber_table.npz"))
    a = ap.parse_args()
    pts = list(range(13))
    tasks = [(e, f) for e in pts for f in range(a.frames)]
    with mp.get_context("fork").Pool(a.procs) as pool:
        res = pool.map(_one, tasks, chunksize=4)
    pt = np.zeros(13, dtype=np.int64)
    pam = np.zeros(13, dtype=np.int64)
    n = np.zeros(13, dtype=np.int64)
    ptm1 = np.zeros(13, dtype=np.int64)
    nm1 = np.zeros(13, dtype=np.int64)
    for e, a_pt, a_pam, c, a_ptm1, cm1 in res:
        pt[e] += a_pt
        pam[e] += a_pam
        n[e] += c
        ptm1[e] += a_ptm1
        nm1[e] += cm1
    # pt_m1_*: the PT detector at timing offset -1, the operating point the shipped example uses
    # for TG (soqpsk_detection.py:184-185); pt_errors / pam_errors are at offset 0.
    np.savez_compressed(a.out, ebn0_db=np.array(pts), pt_errors=pt, pam_errors=pam, compared=n,
                        pt_m1_errors=ptm1, pt_m1_compared=nm1, frames=np.array(a.frames))
    for e in pts:
        print(e, pt[e], pam[e], n[e], ptm1[e], nm1[e])


if __name__ == "__main__":
    main()
