"""Kaiser FIR low-pass front end -- waveforms/filters/lpf.py:13-34 (design, host) plus what the reference
only sketches: applying it to a batch of received frames and choosing the sampling phase."""
from __future__ import annotations

import numpy as np
import torch
from scipy.signal import firwin, kaiserord

from .. import _dev, _ffi
from ..mfbank import filter_bank


def kaiser_fir_lpf(sps: int, f_cutoff: float, width: float | None = None, ripple_db: float = 80.0) -> np.ndarray:
    """Kaiser-window FIR low-pass taps (lpf.py:29-34): kaiserord(ripple_db, width or 1/sps) gives the tap
    count and beta, firwin designs for cutoff = f_cutoff / (sps / 2).  Host float64, set-up work."""
    nyq_rate = sps / 2
    numtaps, beta = kaiserord(ripple_db, width or 1 / sps)
    return firwin(numtaps=numtaps, cutoff=f_cutoff / nyq_rate, window=("kaiser", beta))


def fir_filter(signal: torch.Tensor, taps) -> torch.Tensor:
    """np.convolve(signal, taps, "same") for every row of complex[B, T] (or [T]) on the GPU: the full-rate
    (sps = 1) case of the decimated wf_mf_bank kernel.  Taps real or complex, host."""
    signal = _dev.require_cuda(signal, "signal")
    rows, was_1d = _dev.as_batch(signal)
    taps = np.atleast_2d(np.asarray(taps))
    y = filter_bank(rows, taps, sps=1, first_instant=0, n_calls=rows.shape[1])
    y = y[..., 0]
    return y[0] if was_1d else y


def estimate_timing_offset(y: torch.Tensor, sps: int):
    """Sampling phase of a full-rate matched-filter output y complex64[B, T] (or [T]): the phase n mod sps
    that maximises the mean |y[n]|^2 (wf_timing_metric).  Returns (timing_offset int tensor [B] in the
    reference's sign convention -- the detector uses samples with (n + timing_offset) % sps == 0, so
    timing_offset = -phase -- and the metric float32 [B, sps])."""
    y = _dev.require_cuda(y, "y", torch.complex64)
    rows, was_1d = _dev.as_batch(y)
    B, T = rows.shape
    out = torch.empty((B, sps), dtype=torch.float32, device=rows.device)
    with torch.cuda.device(rows.device):
        _ffi.check(_ffi.lib().wf_timing_metric(_dev.ptr(rows), B, T, T, int(sps), _dev.ptr(out), _dev.stream_ptr(rows.device)))
    off = -torch.argmax(out, dim=1)
    return (off[0] if was_1d else off), (out[0] if was_1d else out)
