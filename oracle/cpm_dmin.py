"""Minimum normalised squared Euclidean distance of a CPM scheme (TEST INFRASTRUCTURE ONLY) --
the textbook external anchor for the two detectors the reference does not have.

Anderson, Aulin & Sundberg, "Digital Phase Modulation" (Plenum, 1986), ch. 3: for an optimum
(MLSE) coherent detector at high Eb/N0

    P_b  ~  C * Q( sqrt( d2_min * Eb/N0 ) ),     d2 = (1 / 2 Eb) * integral |s(t, a) - s(t, a')|^2 dt,

with C of order one.  |s - s'|^2 = 2 - 2 cos(dphi) depends on the two data sequences only through
their DIFFERENCE gamma_k = a_k - a'_k, dphi(t) = 2 pi sum_k h_k gamma_k q(t - kT); for a multi-h
scheme also through which h index meets the first non-zero difference.  The search below uses the
frequency pulses the reference designs (restated in oracle/cpm_link.py and pinned by golden G3) and
enumerates difference sequences of up to `span` symbols that MERGE (accumulated phase difference a
multiple of 2 pi once the pulses have ended).

Known values reproduced (asserted in tests/test_oracle_golden.py): binary 1REC CPFSK h = 0.715 ->
d2_min = 2.43 (the textbook maximum for binary CPFSK).
"""
from __future__ import annotations

from fractions import Fraction
from itertools import product
from math import erfc, sqrt

import numpy as np


def dmin2(pulse, h_list, M, sps, span=5):
    """(d2_min, best difference sequence, first-h index).  h_list: Fractions or floats (cycled per symbol)."""
    g = np.asarray(pulse, dtype=np.float64)
    q = np.concatenate(([0.0], np.cumsum(g))) / sps          # phase pulse samples, q[-1] = 1/2
    K = g.size
    L = -(-K // sps)                                          # pulse length in symbols (ceil)
    hs = [Fraction(h).limit_denominator(64) for h in h_list]
    bps = int(np.log2(M))
    gam_vals = [2 * d for d in range(-(M - 1), M)]            # a - a' with a, a' in {-(M-1), .., M-1} step 2
    n_tot = (span + L) * sps
    best = (np.inf, None, None)
    for start in range(len(hs)):
        hh = [hs[(start + k) % len(hs)] for k in range(span)]
        for gam in product(gam_vals, repeat=span):
            if gam[0] <= 0:                                   # first difference non-zero; sign symmetry halves the search
                continue
            # merge: final phase difference 2 pi * sum h_k gamma_k * q_final = pi * sum h_k gamma_k must be a multiple of 2 pi
            acc = sum(h * gk for h, gk in zip(hh, gam))
            if (acc / 2).denominator != 1:
                continue
            dphi = np.zeros(n_tot)
            for k, gk in enumerate(gam):
                if gk == 0:
                    continue
                idx = np.clip(np.arange(n_tot) + 1 - k * sps, 0, K)
                dphi += 2 * np.pi * float(hh[k]) * gk * q[idx]
            d2 = float(np.sum(2 - 2 * np.cos(dphi))) * bps / (2 * sps)
            if d2 < best[0] - 1e-12:
                best = (d2, gam, start)
    return best


def q_func(x):
    return 0.5 * erfc(x / sqrt(2))


def ebn0_db_for_ber(d2, ber):
    """Eb/N0 (dB) at which Q(sqrt(d2 Eb/N0)) = ber."""
    lo, hi = 0.0, 40.0
    for _ in range(200):
        mid = (lo + hi) / 2
        if q_func(mid) > ber:
            lo = mid
        else:
            hi = mid
    return 10 * np.log10(lo * lo / d2)
