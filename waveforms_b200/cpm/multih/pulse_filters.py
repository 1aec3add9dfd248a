"""ARTM multi-h CPM frequency pulse (host, float64) -- mirrors
waveforms/cpm/multih/pulse_filters.py:7-23: raised cosine over ``length`` symbols (3RC),
h alternating between 4/16 and 5/16."""
from __future__ import annotations

import numpy as np

from ..helpers import normalize_cpm_filter

MULTIH_IRIG_NUMER = np.array([4, 5])
MULTIH_IRIG_DENOM = 16


def freq_pulse_multih_irig(sps: int = 8, length: int = 3) -> np.ndarray:
    t = np.linspace(0, length, num=length * sps + 1)
    return normalize_cpm_filter(sps=sps, g=(1 - np.cos(2 * np.pi * t / length)) / (2 * length))
