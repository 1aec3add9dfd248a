"""Regenerate the reference BER waterfall of BASELINE.md section 2c by running the
REFERENCE ITSELF (read-only at /root/reference): SOQPSK-TG, sps 8, N=4096, Eb/N0 0..12 dB,
--frames per point (committed table: 1024), frame f = PN15 bits [f*4096, ...) mod 32767, noise
PCG64(1000*EbN0_dB + f + salt), all detector rows on the same received frame.  The committed table pools
REPLICATES of the same 1024 frames with independent noise (--salt 0, 1000000, 2000000; --merge): the spread
between replicates is the honest measure of the reference's own sampling error (bit errors come in bursts,
so a binomial interval on one replicate is too narrow).

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
    ebn0, f, salt = args if len(args) == 3 else (*args, 0)
    import make_golden as mg  # imports the reference from /root/reference

    sps = 8
    seq = np.array(mg.PNSequence(15).generate_sequence(), dtype=np.uint8)
    bits = seq[(f * 4096 + np.arange(4096)) % seq.size]
    sigma = float(np.sqrt(sps / (2 * 10 ** (ebn0 / 10))))
    rng = np.random.Generator(np.random.PCG64(seed=1000 * ebn0 + f + salt))
    r = mg.replay_frame(bits, mg.freq_pulse_soqpsk_tg(sps), sps, sigma, rng, {"PT": 0, "PAM": 0, "PT@-1": -1})
    return (ebn0, r["PT"]["bit_errors"], r["PAM"]["bit_errors"], r["PAM"]["compared"],
            r["PT@-1"]["bit_errors"], r["PT@-1"]["compared"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--out", default=os.path.join(HERE, "..", "tests", "golden", "g10_ber_table.npz"))
    ap.add_argument("--salt", type=int, default=0, help="added to every PCG64 seed: an independent replicate on the same frames")
    ap.add_argument("--merge", nargs="+", default=None, help="replicate files to pool into --out (sums + per-replicate counts)")
    a = ap.parse_args()
    if a.merge:
        reps = [np.load(f) for f in a.merge]
        keys = ("pt_errors", "pam_errors", "compared", "pt_m1_errors", "pt_m1_compared")
        out = {k: sum(r[k] for r in reps) for k in keys}
        for k in ("pt_errors", "pam_errors", "pt_m1_errors"):
            out[k + "_by_replicate"] = np.stack([r[k] for r in reps])
        withf = [r for r in reps if "errors_by_frame" in r.files]
        if withf:  # [replicate][point][frame][PT, PAM, PT@-1] of the replicates that recorded it
            out["errors_by_frame"] = np.stack([r["errors_by_frame"] for r in withf])
        np.savez_compressed(a.out, ebn0_db=reps[0]["ebn0_db"], frames=reps[0]["frames"], replicates=np.array(len(reps)),
                            salts=np.array([int(r["salt"]) if "salt" in r.files else 0 for r in reps]), **out)
        for e in range(13):
            print(e, out["pt_errors"][e], out["pam_errors"][e], out["compared"][e], out["pt_m1_errors"][e])
        return
    pts = list(range(13))
    tasks = [(e, f, a.salt) for e in pts for f in range(a.frames)]
    with mp.get_context("fork").Pool(a.procs) as pool:
        res = pool.map(_one, tasks, chunksize=4)
    pt = np.zeros(13, dtype=np.int64)
    pam = np.zeros(13, dtype=np.int64)
    n = np.zeros(13, dtype=np.int64)
    ptm1 = np.zeros(13, dtype=np.int64)
    nm1 = np.zeros(13, dtype=np.int64)
    # per-frame error counts (task order = point-major, frame-minor): bit errors come in bursts, so the
    # honest variance of a point's error count is the between-frame one, not the binomial one
    by_frame = np.array([[r[1], r[2], r[4]] for r in res], dtype=np.int32).reshape(13, a.frames, 3)
    for e, a_pt, a_pam, c, a_ptm1, cm1 in res:
        pt[e] += a_pt
        pam[e] += a_pam
        n[e] += c
        ptm1[e] += a_ptm1
        nm1[e] += cm1
    # pt_m1_*: the PT detector at timing offset -1, the operating point the shipped example uses
    # for TG (soqpsk_detection.py:184-185); pt_errors / pam_errors are at offset 0.
    np.savez_compressed(a.out, ebn0_db=np.array(pts), pt_errors=pt, pam_errors=pam, compared=n,
                        pt_m1_errors=ptm1, pt_m1_compared=nm1, frames=np.array(a.frames), salt=np.array(a.salt),
                        errors_by_frame=by_frame)  # [point][frame][PT, PAM, PT@-1]
    for e in pts:
        print(e, pt[e], pam[e], n[e], ptm1[e], nm1[e])


if __name__ == "__main__":
    main()
