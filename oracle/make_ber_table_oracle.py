"""Extend the reference BER table with REFERENCE-EQUIVALENT replicates (test infrastructure only).

One replicate of oracle/make_ber_table.py -- the reference itself, pure Python -- costs about 100 minutes on four
cores.  The oracle chain (oracle/cpm_link.py) reproduces the reference's error count of every frame exactly when it is
fed the same PN frame and the same PCG64 seed (tests/test_oracle_golden.py::
test_oracle_link_reproduces_reference_ber_table_frames; this script's --check reproduces whole replicates of G10 cell by
cell: done for salts 0 and 1e6), and with its detector vectorised over frames one replicate takes a few minutes.  This script therefore runs the
SAME experiment (frames 0..1023 x 13 points, PCG64(1000 * EbN0 + frame + salt), PT / PAM at offset 0 and PT at offset
-1 on the same received frame) for further salts and writes tests/golden/g10b_ber_table_oracle.npz: per replicate and
pooled error counts, flagged source = "oracle".

    python oracle/make_ber_table_oracle.py --salts 5000000 6000000 ... [--procs 8] [--check 0 1000000]
"""
from __future__ import annotations

import argparse
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import cpm_link as o  # noqa: E402

N_BITS, SPS, FRAMES = 4096, 8, 1024
_PULSE = None


def _chunk(args):
    """16 frames of one (point, salt): returns int32[16, 3] bit errors (PT, PAM, PT@-1) and the compared counts."""
    global _PULSE
    ebn0, f0, nf, salt = args
    if _PULSE is None:
        _PULSE = o.pulse_soqpsk_tg(SPS)
    pulse = _PULSE
    cols = o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"]
    sigma = float(np.sqrt(SPS / (2 * 10 ** (ebn0 / 10))))
    zs = {"PT": [], "PAM": [], "PT@-1": []}
    bits_all, syms_all = [], []
    for f in range(f0, f0 + nf):
        bits = o.pn_frame_bits(15, f, N_BITS)
        syms = o.fsm_encode(cols, bits)
        s = o.cpm_modulate_closed_form(syms, 0.25, pulse, SPS)
        rng = np.random.Generator(np.random.PCG64(seed=1000 * ebn0 + f + salt))
        rx = s * np.exp(-1j * np.pi / 4) + o.complex_awgn(sigma, s.size, rng)
        pt, pam = o.pt_bank(rx, pulse, 0.25, SPS), o.pam_bank(rx, pulse, 0.25, SPS)
        i0, im1 = o.symbol_instants(rx.size, SPS, 0, 2), o.symbol_instants(rx.size, SPS, -1, 2)
        zs["PT"].append(pt[:, i0].T)
        zs["PAM"].append(pam[:, i0].T)
        zs["PT@-1"].append(pt[:, im1].T)
        bits_all.append(bits)
        syms_all.append(syms)
    out = np.zeros((nf, 3), dtype=np.int32)
    cmp_ = np.zeros(3, dtype=np.int64)
    for c, key in enumerate(("PT", "PAM", "PT@-1")):
        db, ds = o.detect_soqpsk_fast(np.stack(zs[key]), True)
        for k in range(nf):
            e, _se, n = o.count_errors(db[k], ds[k], bits_all[k], syms_all[k], 2)
            out[k, c] = e
            cmp_[c] += n
    return ebn0, f0, out, cmp_


def run_salt(salt, pool):
    tasks = [(e, f0, 16, salt) for e in range(13) for f0 in range(0, FRAMES, 16)]
    by_frame = np.zeros((13, FRAMES, 3), dtype=np.int32)
    compared = np.zeros((13, 3), dtype=np.int64)
    for e, f0, out, c in pool.imap_unordered(_chunk, tasks, chunksize=2):
        by_frame[e, f0:f0 + out.shape[0]] = out
        compared[e] += c
    return by_frame, compared


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--salts", type=int, nargs="*", default=[])
    ap.add_argument("--check", type=int, nargs="*", default=[], help="salts of G10 replicates to reproduce cell by cell first")
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--out", default=os.path.join(HERE, "..", "tests", "golden", "g10b_ber_table_oracle.npz"))
    a = ap.parse_args()
    g10 = np.load(os.path.join(HERE, "..", "tests", "golden", "g10_ber_table.npz"))
    with mp.get_context("fork").Pool(a.procs) as pool:
        checked = []
        for salt in a.check:
            t0 = time.time()
            by_frame, compared = run_salt(salt, pool)
            r = [int(x) for x in g10["salts"]].index(salt)
            tot = by_frame.sum(axis=1)
            for c, key in enumerate(("pt_errors", "pam_errors", "pt_m1_errors")):
                assert np.array_equal(tot[:, c], g10[key + "_by_replicate"][r]), (salt, key, tot[:, c], g10[key + "_by_replicate"][r])
            print(f"check salt {salt}: all 39 (point, detector) totals equal the reference's replicate ({time.time() - t0:.0f} s)", flush=True)
            checked.append(salt)
        reps, cmps = [], []
        for salt in a.salts:
            t0 = time.time()
            by_frame, compared = run_salt(salt, pool)
            reps.append(by_frame.sum(axis=1))
            cmps.append(compared)
            print(f"salt {salt}: PT {reps[-1][4, 0]} / PAM {reps[-1][4, 1]} errors at 4 dB ({time.time() - t0:.0f} s)", flush=True)
    if reps:
        reps, cmps = np.stack(reps), np.stack(cmps)          # [salt][point][PT, PAM, PT@-1]
        np.savez_compressed(a.out, source=np.array("oracle"), ebn0_db=np.arange(13), frames=np.array(FRAMES),
                            salts=np.array(a.salts), checked_against_g10_salts=np.array(checked, dtype=np.int64),
                            errors_by_replicate=reps.astype(np.int64), compared_by_replicate=cmps,
                            pt_errors=reps[:, :, 0].sum(0), pam_errors=reps[:, :, 1].sum(0), pt_m1_errors=reps[:, :, 2].sum(0),
                            compared=cmps[:, :, 0].sum(0), pt_m1_compared=cmps[:, :, 2].sum(0))
        print("wrote", a.out)


if __name__ == "__main__":
    main()
