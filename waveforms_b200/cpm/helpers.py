"""Host-side pulse helpers -- mirrors waveforms/cpm/helpers.py:5-19."""
from __future__ import annotations

import numpy as np


def normalize_cpm_filter(sps: int, g: np.ndarray) -> np.ndarray:
    """Scale a frequency pulse so that sum(g) = sps/2, i.e. its integral is 1/2."""
    g = np.asarray(g, dtype=np.float64)
    return g * (sps / (2.0 * g.sum()))
