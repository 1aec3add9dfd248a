"""Float64 numpy restatement of the ARRAYS behind the reference's diagnostic figures
(TEST INFRASTRUCTURE ONLY) -- SURVEY 8(f)3/4.  The reference's waveforms/viz/*.py import matplotlib at
module level (absent here; figure output is unavailable on this box), so the lines that compute what is
plotted are restated, each citing the reference file:line it follows.

``psd`` restates a THIRD-PARTY algorithm: viz/psd.py:35-40 calls matplotlib's ``Axes.psd`` =
``matplotlib.mlab.psd`` (pinned 3.7.5 / 3.9.4 / 3.10.1 by the reference's uv.lock, not installed in this
image).  Its published algorithm (mlab._spectral_helper, mode "psd"): split x into NFFT-sample segments
with ``noverlap`` = 0 (a trailing partial segment is dropped), no detrending, multiply by the symmetric
Hann window np.hanning(NFFT), FFT, |X|^2, average over segments, divide by (sum |window|)^2 when
scale_by_freq is False; complex input -> two-sided, returned fft-shifted with
freqs = fftshift(fftfreq(NFFT, 1/Fs)).  PARITY UNPINNED against matplotlib itself (it cannot run here).
"""
from __future__ import annotations

import numpy as np


def psd(signal, sps, bps=1, nfft=1024):
    """viz/psd.py:35-40: Axes.psd(signal * sqrt(bps), NFFT=nfft, Fs=sps/bps, scale_by_freq=False).
    Returns (freqs, Pxx) -- linear power; the figure shows 10 log10(Pxx)."""
    x = np.asarray(signal, dtype=np.complex128) * np.sqrt(bps)
    n_seg = x.size // nfft
    win = np.hanning(nfft)
    seg = x[: n_seg * nfft].reshape(n_seg, nfft) * win
    pxx = np.mean(np.abs(np.fft.fft(seg, axis=1)) ** 2, axis=0) / np.abs(win).sum() ** 2
    freqs = np.fft.fftshift(np.fft.fftfreq(nfft, d=bps / sps))
    return freqs, np.fft.fftshift(pxx)


def eye_traces(time, signal, sps=8, modulo=4, t_offset=0.0):
    """viz/eye.py:39-56: for i in range((time.size - 1) // (sps * modulo)) the trace
    signal[i*span : i*span + span + 1] against time[idx : idx + span + 1] - time[idx] + t_offset."""
    span = sps * modulo
    n = (np.asarray(time).size - 1) // span
    t = np.stack([time[i * span: i * span + span + 1] - time[i * span] + t_offset for i in range(n)]) if n else np.zeros((0, span + 1))
    tr = np.stack([signal[i * span: i * span + span + 1] for i in range(n)]) if n else np.zeros((0, span + 1), complex)
    return t, tr


def constellation(signal, n=1024):
    """viz/constellation.py:30-36: the first n samples, in-phase against quadrature."""
    s = np.asarray(signal)[:n]
    return s.real, s.imag


def phase_tree(signal, sps, off=None, modulo=4):
    """viz/tree.py:64-71: per chunk of modulo*sps samples, np.unwrap(np.angle(.)) minus `off` (or the
    chunk's first value); x axis np.linspace(0, modulo, modulo*sps, endpoint=False)."""
    phase = np.angle(np.asarray(signal))
    span = sps * modulo
    t = np.linspace(0, modulo, span, endpoint=False)
    rows = []
    for chunk in range(phase.size // span):
        sl = np.unwrap(phase[chunk * span: (chunk + 1) * span])
        rows.append(sl - (off if off is not None else sl[0]))
    return t, (np.stack(rows) if rows else np.zeros((0, span)))


def kaiser_fir_lpf(sps, f_cutoff, width=None, ripple_db=80.0):
    """filters/lpf.py:13-34 (scipy.signal.kaiserord + firwin; scipy is installed here)."""
    from scipy.signal import firwin, kaiserord

    numtaps, beta = kaiserord(ripple_db, width or 1 / sps)
    return firwin(numtaps=numtaps, cutoff=f_cutoff / (sps / 2), window=("kaiser", beta))


def timing_metric(y, sps):
    """|mean_i (-1)^i y[ph + sps i]^2| per sampling phase ph (the build's replacement for the hard-coded timing
    offsets of examples/soqpsk_detection.py:183-192; no reference implementation).  For offset-QPSK-like CPM the
    matched-filter output alternates between the I and the Q axis from bit to bit at the right sampling phase, so
    y^2 alternates in sign there; |y|^2 itself is flat (constant envelope) and carries no timing information."""
    y = np.asarray(y)
    out = []
    for ph in range(sps):
        v = y[ph::sps] ** 2
        sign = 1.0 - 2.0 * (np.arange(v.size) & 1)
        out.append(abs(np.sum(sign * v)) / max(v.size, 1))
    return np.array(out)
