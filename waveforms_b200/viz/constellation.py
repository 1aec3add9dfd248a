"""Constellation samples -- waveforms/viz/constellation.py:15-44 without the figure."""
from __future__ import annotations

import torch

from .. import _dev


def constellation(signal: torch.Tensor, n: int = 1024, color=None, axis=None):
    """(in-phase, quadrature) of the first n samples (constellation.py:30-36): views of the signal, no copy."""
    signal = _dev.require_cuda(signal, "signal")
    s = torch.view_as_real(signal[..., :n])
    return s[..., 0], s[..., 1]
