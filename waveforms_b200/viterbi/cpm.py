"""Generic pulse-truncated coherent CPM trellis detection on the GPU (rows A11/A12).

The reference has no detector for PCM/FM or ARTM multi-h CPM (waveforms.viterbi exports
only SOQPSKTrellisDetector); the waveform parameters come from
examples/irig_comparison.py:50-72.  ``CPMTrellisDetector`` generalises the reference's
pulse-truncation bank (examples/soqpsk_detection.py:134-156) and its
Re{e^{-j theta_state} z} branch metric (viterbi/algorithm.py:57-63) to Lp >= 1 symbols of
memory, M-ary alphabets and cycling modulation indices h_k = h_num[k % n_h] / p:

    bank    M^Lp complex sps-tap correlators per step parity (host design, wf_mf_bank)
    trellis 2p * M^(Lp-1) states, register-exchange survivors      (wf_viterbi_cpm)

Model: with Q the cumulative frequency pulse, off = (K-1)//2 and Qn[j] = Q[clip(j,0,K)]/sps the
modulator's phase is 2 pi sum_k h_k a_k Qn[n + off + 1 - (k+1) sps]; Qn is treated as 0 before
index a = ((K-1) - Lp sps)//2 and as complete after a + Lp sps.  Step i covers samples
c_i + u, u in [0, sps), c_i = (i+1) sps - off - 1 + a.  Parity unpinned.
"""
from __future__ import annotations

from itertools import product

import numpy as np
import torch

from .. import _dev, _ffi
from ..mfbank import filter_bank


class CPMTrellisDetector:
    def __init__(self, pulse_filter, h_num, h_denom: int, alphabet_size: int, sps: int = 8, memory: int = 2,
                 depth: int = 0) -> None:
        self.g = np.asarray(pulse_filter, dtype=np.float64)
        self.h_num = np.atleast_1d(np.asarray(h_num, dtype=np.int32)).copy()
        self.p, self.M, self.sps, self.Lp = int(h_denom), int(alphabet_size), int(sps), int(memory)
        self.depth = int(depth)
        K = self.g.size
        self.off = (K - 1) // 2
        self.a = max(0, ((K - 1) - self.Lp * self.sps) // 2)
        self.first_instant = self.sps - self.off - 1 + self.a
        self.n_h = int(self.h_num.size)
        self.n_filt = self.M ** self.Lp
        self.n_states = 2 * self.p * self.M ** (self.Lp - 1)
        self.taps = self._design_taps()

    def _design_taps(self) -> np.ndarray:
        """complex[n_h * M^Lp, sps] in the convolution form wf_mf_bank takes (center = sps-1):
        the conjugated, time-reversed hypothesis waveforms exp(j 2 pi sum_l h_{i-l} a_l Qn[a + l sps + u])."""
        Qn = np.concatenate(([0.0], np.cumsum(self.g))) / self.sps
        u = np.arange(self.sps)
        taps = np.zeros((self.n_h * self.n_filt, self.sps), dtype=np.complex128)
        for par in range(self.n_h):
            for f, hyp in enumerate(product(range(self.M), repeat=self.Lp)):  # hyp[0] = newest symbol
                turns = np.zeros(self.sps)
                for lag, v in enumerate(hyp):
                    h = self.h_num[(par - lag) % self.n_h] / self.p
                    turns += h * (2 * v - (self.M - 1)) * Qn[np.clip(self.a + lag * self.sps + u, 0, self.g.size)]
                taps[par * self.n_filt + f] = np.exp(-2j * np.pi * turns)[::-1]
        return taps

    def bank(self, rx: torch.Tensor, n_steps: int, n_samples: int | None = None) -> torch.Tensor:
        """complex[B, n_steps, n_h * M^Lp] hypothesis correlations of rx complex[B, T]."""
        return filter_bank(rx, self.taps, sps=self.sps, first_instant=self.first_instant, n_calls=n_steps,
                           center=self.sps - 1, n_samples=n_samples)

    def supports_fused(self, rx_dtype, n_samples: int, ld_rx: int, n_steps: int) -> bool:
        """True when wf_cpm_detect has a fused bank+detector kernel for this trellis and signal shape."""
        if rx_dtype not in _dev.COMPLEX:
            return False
        tr, ti = _dev.host_f64(self.taps.real.ravel()), _dev.host_f64(self.taps.imag.ravel())
        return bool(_ffi.lib().wf_cpm_detect_supported(_dev.COMPLEX[rx_dtype], int(n_samples), int(ld_rx), _dev.f64_ptr(tr),
                                                       _dev.f64_ptr(ti), self.sps, self.first_instant, int(n_steps), self.M,
                                                       self.Lp, self.p, self.h_num.ctypes.data, self.n_h, self.depth))

    def detect_rx(self, rx: torch.Tensor, n_steps: int, *, n_samples: int | None = None,
                  tx_symbols: torch.Tensor | None = None, skip_head: int = 4, skip_tail: int = 4,
                  point: torch.Tensor | None = None, counters: torch.Tensor | None = None, return_decisions: bool = True,
                  tune: int = 0):
        """Detection straight from the received samples rx complex64[B, ld >= T] (wf_cpm_detect): the
        hypothesis bank runs inside the detector kernel.  Same outputs as ``detect(bank(rx, n_steps))``;
        raises _ffi.WfError for shapes without a fused kernel (see ``supports_fused``)."""
        rx = _dev.require_cuda(rx, "rx")
        if rx.dtype not in _dev.COMPLEX or rx.dim() != 2:
            raise TypeError("rx must be complex[B, T]")
        B, ld = rx.shape
        T = ld if n_samples is None else int(n_samples)
        dec = torch.empty((B, n_steps), dtype=torch.int8, device=rx.device) if return_decisions else None
        S = 0
        if tx_symbols is not None:
            tx_symbols = _dev.require_cuda(tx_symbols, "tx_symbols", torch.int8)
            S = tx_symbols.shape[1]
            if counters is None:
                raise ValueError("tx_symbols given without counters")
        if counters is not None:
            counters = _dev.require_cuda(counters, "counters", torch.int64)
        if point is not None:
            point = _dev.require_cuda(point, "point", torch.int32)
        tr, ti = _dev.host_f64(self.taps.real.ravel()), _dev.host_f64(self.taps.imag.ravel())
        with torch.cuda.device(rx.device):
            _ffi.check(_ffi.lib().wf_cpm_detect(_dev.ptr(rx), _dev.COMPLEX[rx.dtype], B, T, ld, _dev.f64_ptr(tr), _dev.f64_ptr(ti),
                                                self.sps, self.first_instant, int(n_steps), self.M, self.Lp, self.p,
                                                self.h_num.ctypes.data, self.n_h, self.depth, _dev.ptr(dec),
                                                _dev.ptr(tx_symbols), S, int(skip_head), int(skip_tail), _dev.ptr(point),
                                                _dev.ptr(counters), int(tune), _dev.stream_ptr(rx.device)))
        return dec

    def detect(self, z: torch.Tensor, *, tx_symbols: torch.Tensor | None = None, skip_head: int = 4, skip_tail: int = 4,
               point: torch.Tensor | None = None, counters: torch.Tensor | None = None, return_decisions: bool = True):
        """z complex[B, n_steps, n_h * M^Lp] -> detected symbols int8[B, n_steps] (alphabet values).
        With ``counters`` int64[P, 3] and ``tx_symbols`` int8[B, S]: (bit errors, symbol errors, compared
        BITS = log2(M) per symbol) over symbols [skip_head, min(S, n_steps) - skip_tail) are added at row
        point[b], so counters[:, 0] / counters[:, 2] is a BER for every alphabet."""
        z = _dev.require_cuda(z, "z")
        if z.dtype not in _dev.COMPLEX or z.dim() != 3 or z.shape[2] != self.n_h * self.n_filt:
            raise TypeError(f"z must be complex[B, steps, {self.n_h * self.n_filt}]")
        B, n_steps, _ = z.shape
        dec = torch.empty((B, n_steps), dtype=torch.int8, device=z.device) if return_decisions else None
        S = 0
        if tx_symbols is not None:
            tx_symbols = _dev.require_cuda(tx_symbols, "tx_symbols", torch.int8)
            S = tx_symbols.shape[1]
            if counters is None:
                raise ValueError("tx_symbols given without counters")
        if counters is not None:
            counters = _dev.require_cuda(counters, "counters", torch.int64)
        if point is not None:
            point = _dev.require_cuda(point, "point", torch.int32)
        with torch.cuda.device(z.device):
            _ffi.check(_ffi.lib().wf_viterbi_cpm(_dev.ptr(z), _dev.COMPLEX[z.dtype], B, n_steps, n_steps, self.M, self.Lp,
                                                 self.p, self.h_num.ctypes.data, self.n_h, self.depth, _dev.ptr(dec),
                                                 _dev.ptr(tx_symbols), S, int(skip_head), int(skip_tail),
                                                 _dev.ptr(point), _dev.ptr(counters), _dev.stream_ptr(z.device)))
        return dec
