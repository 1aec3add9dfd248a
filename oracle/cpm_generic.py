"""Float64 numpy statement of the generic pulse-truncated coherent CPM trellis detector
(TEST INFRASTRUCTURE ONLY) -- rows A11/A12 of SURVEY section 8(a).

PARITY UNPINNED: the reference has NO detector for PCM/FM or ARTM multi-h CPM
(waveforms.viterbi exports only SOQPSKTrellisDetector).  This is the builder's own design,
generalising the reference's pulse-truncation idea (examples/soqpsk_detection.py:134-156:
filters exp(-2j pi h alpha q(t)) over the central symbol of the phase pulse, metric
Re{e^{-j theta_state} z}) to L' >= 1 symbols of memory, M-ary alphabets and cycling
modulation indices (waveform parameters from examples/irig_comparison.py:50-72).
It is validated by noiseless round trips, by the modulators (which ARE pinned, G4) and by
monotone BER curves; the CUDA kernel (csrc/cpmdet.cu) is tested against it decision by decision.

Model.  With Q the cumulative frequency pulse, off = (K-1)//2 and
Qn[j] = Q[clip(j,0,K)]/sps, the modulator's phase is 2 pi sum_k h_k a_k Qn[n + off + 1 - (k+1) sps].
Truncation to Lp symbols: Qn is taken as 0 before index a = ((K-1) - Lp sps)//2 and as its final
value after a + Lp sps.  Detection step i covers samples c_i + u, u in [0, sps),
c_i = (i+1) sps - off - 1 + a, where symbols i, i-1, .., i-Lp+1 are in transit and older ones
contribute the phase state theta_i = pi/p sum_{k <= i-Lp} K_k a_k  (h_k = K_k / p).
"""
from __future__ import annotations

from itertools import product

import numpy as np


class CpmTrellis:
    def __init__(self, pulse, h_num, p, M, sps, Lp):
        self.g = np.asarray(pulse, dtype=np.float64)
        self.h_num = [int(x) for x in np.atleast_1d(h_num)]
        self.p, self.M, self.sps, self.Lp = int(p), int(M), int(sps), int(Lp)
        K = self.g.size
        self.off = (K - 1) // 2
        self.a = max(0, ((K - 1) - self.Lp * self.sps) // 2)
        self.first_instant = self.sps - self.off - 1 + self.a  # c_0
        self.n_h = len(self.h_num)
        self.alphabet = [2 * v - (self.M - 1) for v in range(self.M)]
        self.n_filt = self.M ** self.Lp
        self.n_hist = self.M ** (self.Lp - 1)
        self.n_states = 2 * self.p * self.n_hist
        Q = np.concatenate(([0.0], np.cumsum(self.g)))
        self.Qn = Q / self.sps
        self.full = float(self.Qn[-1])  # 1/2 for a normalised pulse

    # hypothesis index f = sum_l v_l M^(Lp-1-l), v_0 = newest symbol (a_l = alphabet[v_l])
    def templates(self):
        """complex[n_h, n_filt, sps]: exp(+j 2 pi sum_l h_{i-l} a_l Qn[a + l sps + u]) for step parity i % n_h."""
        out = np.zeros((self.n_h, self.n_filt, self.sps), dtype=np.complex128)
        u = np.arange(self.sps)
        for par in range(self.n_h):
            for f, vs in enumerate(product(range(self.M), repeat=self.Lp)):
                ph = np.zeros(self.sps)
                for l, v in enumerate(vs):
                    h = self.h_num[(par - l) % self.n_h] / self.p
                    idx = np.clip(self.a + l * self.sps + u, 0, self.g.size)
                    ph += h * self.alphabet[v] * self.Qn[idx]
                out[par, f] = np.exp(2j * np.pi * ph)
        return out

    def bank_taps(self):
        """taps complex[n_h * n_filt, sps] for the convolution-form bank
        y[i] = sum_t rx[c_i + (sps-1) - t] taps[t]  (center = sps-1): conj(template) reversed."""
        t = self.templates().reshape(self.n_h * self.n_filt, self.sps)
        return np.conj(t[:, ::-1])

    def bank(self, rx, n_steps):
        """z complex[n_steps, n_h * n_filt] by direct correlation (rx = 0 outside the frame)."""
        tpl = np.conj(self.templates().reshape(self.n_h * self.n_filt, self.sps))
        z = np.zeros((n_steps, tpl.shape[0]), dtype=np.complex128)
        for i in range(n_steps):
            c = self.first_instant + i * self.sps
            seg = np.zeros(self.sps, dtype=np.complex128)
            lo, hi = max(c, 0), min(c + self.sps, rx.size)
            if hi > lo:
                seg[lo - c:hi - c] = rx[lo:hi]
            z[i] = tpl @ seg
        return z

    def default_depth(self):
        bits = max(1, (self.M - 1).bit_length())
        return max(1, min(16, 32 // bits))

    def start_parity(self):
        """Parity of the phase-state index v the detector may start from, or None when every v is reachable.

        Every transition of step i moves v by K_i a with a odd, so it changes the parity of v by K_i mod 2 -- the
        same for all branches of the step: the trellis is bipartite and at any time only the p phase states of ONE
        parity class are reachable (the textbook count: p states for even K, 2 p for odd K, of which p at a time).
        The frame starts with no symbols in the air, i.e. v = 0 once the Lp - 1 phantom symbols of the first steps
        have left the window; walking those steps backwards gives the parity class at the start."""
        if all(k % 2 == 0 for k in self.h_num):
            return None                      # even K only: parity never changes; odd v are the same dead class forever
        par = 0
        for i in range(self.Lp - 2, -1, -1):  # steps whose leaving symbol index i - Lp + 1 is negative (phantom)
            par = (par - self.h_num[(i - self.Lp + 1) % self.n_h]) % 2
        return par

    def live_parity(self, i):
        """Parity class of v AFTER step i (start_parity() propagated through steps 0..i)."""
        par = self.start_parity()
        if par is None:
            return None
        for j in range(i + 1):
            par = (par + self.h_num[(j - self.Lp + 1) % self.n_h]) % 2
        return par

    def detect(self, z, depth=None, return_margin=False, reachable=True):
        """Viterbi detection maximising sum Re{e^{-j pi v/p} z}.  z complex[n_steps, n_h*n_filt].
        Returns symbol INDICES (0..M-1) uint8[n_steps].  The history part of the start state is unknown (all equal);
        the phase part starts in the reachable parity class (``reachable``: states of the other class start at
        -inf and stay there -- see start_parity; False = all 2 p phase states start equal, the round-1 rule).

        Decision rule -- BLOCK RELEASE with survivor memory 2 D (D = ``depth``, default min(16, 32 / log2 M)):
        the add-compare-select of every step stores, per state, which predecessor won; whenever
        (t + 1) is a multiple of D and t + 1 >= 2 D the survivor of the best state at step t (largest
        metric, lowest index on ties) is traced back 2 D steps and the symbols of steps
        [t - 2D + 1, t - D] are released -- every decision is taken with at least D steps of look-ahead;
        after the last step the survivor of the best final state releases everything still held.
        (A register-exchange detector with 2 D-symbol survivor words released D at a time is the same rule.)

        ``return_margin``: also return float64[n_steps, 2] = per step (smallest gap between the best and
        the runner-up candidate over all states and between the two largest new path metrics, largest
        |candidate|).  The decision for symbol k is released at a step < k + 2 D, so a float32 detector may
        differ at k only if some step in [k, k + 2 D) has a margin at float32 resolution."""
        M, Lp, p2 = self.M, self.Lp, 2 * self.p
        D = int(depth or self.default_depth())
        n_steps = z.shape[0]
        NS, NH = self.n_states, self.n_hist
        pm = np.zeros(NS)
        sp0 = self.start_parity() if reachable else None
        if sp0 is not None:
            pm[(np.arange(NS) // NH) % 2 != sp0] = -np.inf
        rot = np.exp(-1j * np.pi * np.arange(p2) / self.p)
        out = np.zeros(n_steps, dtype=np.uint8)
        marg = np.zeros((n_steps, 2))
        pred_of = np.zeros((n_steps, NS), dtype=np.int32)   # winning predecessor state per step and state
        sym_of = np.zeros((n_steps, NS), dtype=np.uint8)    # symbol index of step i on that branch
        sp = np.arange(NS)
        vnew, hnew = sp // NH, sp % NH          # next state = (v', [a_0 .. a_{Lp-2}])
        released = 0

        def release(t, upto):
            """trace the best state at step t back to step `released`; emit steps [released, upto)."""
            s_ = int(np.argmax(pm))
            for tau in range(t, released - 1, -1):
                if tau < upto:
                    out[tau] = sym_of[tau, s_]
                s_ = int(pred_of[tau, s_])

        for i in range(n_steps):
            par = i % self.n_h
            K_leave = self.h_num[(i - Lp + 1) % self.n_h]
            zi = z[i, par * self.n_filt:(par + 1) * self.n_filt]
            best = np.full(NS, -np.inf)
            second = np.full(NS, -np.inf)
            arg = np.zeros(NS, dtype=np.int64)
            argx = np.zeros(NS, dtype=np.int64)
            amax = 0.0
            for x in range(M):                   # x = index of the symbol leaving the window
                vpred = (vnew - K_leave * self.alphabet[x]) % p2
                if Lp == 1:
                    pred, f = vpred, np.full(NS, x)
                else:
                    pred = vpred * NH + (hnew % (NH // M)) * M + x
                    f = hnew * M + x
                cand = pm[pred] + np.real(rot[vpred] * zi[f])
                take = cand > best               # strict: the lowest x wins ties
                second = np.where(take, best, np.maximum(second, cand))
                amax = max(amax, float(np.max(np.abs(cand[np.isfinite(cand)]))) if np.any(np.isfinite(cand)) else 0.0)
                best = np.where(take, cand, best)
                arg = np.where(take, pred, arg)
                argx = np.where(take, x, argx)
            pred_of[i] = arg
            sym_of[i] = argx if Lp == 1 else hnew // (NH // M)   # symbol index of step i
            pm = best - best.max()
            if return_margin:
                live = np.isfinite(best)
                top = np.sort(best[live])
                gaps = best[live] - second[live]          # second = -inf where a live state has one live candidate only
                gaps = gaps[np.isfinite(gaps)]
                marg[i] = (min(float(np.min(gaps)) if gaps.size else np.inf, float(top[-1] - top[-2])), amax)
            if (i + 1) % D == 0 and i + 1 >= 2 * D:
                release(i, i + 1 - D)
                released = i + 1 - D
        if n_steps:
            release(n_steps - 1, n_steps)
        if return_margin:
            return out, marg
        return out
