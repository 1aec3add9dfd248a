"""PCM/FM frequency pulse (host, float64) -- mirrors waveforms/cpm/pcmfm/pulse_filters.py:8-25:
an NRZ rectangle of one symbol convolved with the impulse response of the
pre-modulation Bessel low-pass (magnitude-normalised prototype), scaled to sum sps/2."""
from __future__ import annotations

import numpy as np
from scipy.signal import besselap, impulse

from ..helpers import normalize_cpm_filter


def freq_pulse_pcmfm(sps: int = 8, order: int = 4) -> np.ndarray:
    span = 3  # symbols covered by the pulse
    n_filter = (span - 1) * sps + 1
    t = np.linspace(0, span * 2 / 0.7, num=n_filter, endpoint=False)
    _, h = impulse(besselap(order, norm="mag"), T=t)
    rect = np.full(sps, 1.0 / (2 * sps))
    return normalize_cpm_filter(sps=sps, g=np.convolve(rect, h, mode="full"))
