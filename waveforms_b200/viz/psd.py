"""Welch power spectral density -- waveforms/viz/psd.py:16-54 without the figure."""
from __future__ import annotations

import numpy as np
import torch

from .. import _dev, _ffi


def plot_power_spectral_density(signal: torch.Tensor, sps: int, bps: int = 1, nfft: int = 1024, axis=None):
    """What ``axis.psd(signal * sqrt(bps), NFFT=nfft, Fs=sps / bps, scale_by_freq=False)`` plots
    (matplotlib.mlab.psd: Hann window, no overlap, two-sided, fft-shifted).

    signal: complex64 CUDA tensor [T] or [B, T].  Returns (freqs float64 numpy [nfft] in units of the bit
    rate, Pxx float32 CUDA tensor [nfft] or [B, nfft], linear power; the figure shows 10 log10(Pxx) dBc)."""
    signal = _dev.require_cuda(signal, "signal", torch.complex64)
    rows, was_1d = _dev.as_batch(signal)
    B, T = rows.shape
    if T < nfft:
        raise ValueError(f"signal shorter ({T}) than one segment of nfft = {nfft} samples")
    out = torch.empty((B, nfft), dtype=torch.float32, device=rows.device)
    with torch.cuda.device(rows.device):
        _ffi.check(_ffi.lib().wf_welch_psd(_dev.ptr(rows), B, T, T, int(nfft), float(np.sqrt(bps)), _dev.ptr(out),
                                           _dev.stream_ptr(rows.device)))
    freqs = np.fft.fftshift(np.fft.fftfreq(nfft, d=bps / sps))
    return freqs, (out[0] if was_1d else out)
