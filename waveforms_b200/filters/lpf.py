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
    """Sampling phase of a full-rate matched-filter output y complex64[B, T] (or [T]) of an offset-QPSK-like
    CPM signal (SOQPSK): the phase ph = n mod sps that maximises |mean_i (-1)^i y[ph + sps i]^2| (wf_timing_metric;
    the envelope of a constant-envelope signal is flat, the sign alternation of y^2 between I and Q bits is not).
    Returns (timing_offset int tensor [B], metric float32 [B, sps]).  timing_offset is in the reference's sign
    convention -- the detector uses samples with (n + timing_offset) % sps == 0 -- as the representative of -ph
    nearest to zero, in [-sps/2, sps/2): the hard-coded values of examples/soqpsk_detection.py:183-192 (0, -1, -2)
    sit on the late shoulder of the same peak (the even-length 'same' convolution puts it half a sample before
    n = 0).  A POSITIVE value means the first usable instant is sps - offset, one bit period into the frame:
    the driver then loses the frame's first decision (skip one more in the tally)."""
    y = _dev.require_cuda(y, "y", torch.complex64)
    rows, was_1d = _dev.as_batch(y)
    B, T = rows.shape
    out = torch.empty((B, sps), dtype=torch.float32, device=rows.device)
    with torch.cuda.device(rows.device):
        _ffi.check(_ffi.lib().wf_timing_metric(_dev.ptr(rows), B, T, T, int(sps), _dev.ptr(out), _dev.stream_ptr(rows.device)))
    ph = torch.argmax(out, dim=1)
    off = torch.where(ph > sps // 2, sps - ph, -ph)
    return (off[0] if was_1d else off), (out[0] if was_1d else out)
