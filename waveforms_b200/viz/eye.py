"""Eye-diagram traces -- waveforms/viz/eye.py:15-65 without the figure."""
from __future__ import annotations

import torch

from .. import _dev, _ffi


def plot_eye_diagram(time: torch.Tensor, signal: torch.Tensor, sps: int = 8, modulo: int = 4, t_offset: float = 0,
                     color=None, axes=None):
    """The traces the reference overlays (eye.py:39-56): trace i = signal[i*span : i*span + span + 1],
    span = sps * modulo, for i < (time.size - 1) // span, against (time[idx : idx+span+1] - time[idx]) + t_offset.

    signal complex64 CUDA [T] or [B, T]; time float64 CUDA [T] (as returned by cpm_modulate).
    Returns (t float64 [span+1], traces complex64 [n_traces, span+1] or [B, n_traces, span+1]); the
    in-phase eye is traces.real, the quadrature eye traces.imag."""
    signal = _dev.require_cuda(signal, "signal", torch.complex64)
    rows, was_1d = _dev.as_batch(signal)
    B, T = rows.shape
    span = int(sps) * int(modulo)
    n_tr = (int(time.shape[-1]) - 1) // span
    out = torch.empty((B, n_tr, span + 1), dtype=torch.complex64, device=rows.device)
    with torch.cuda.device(rows.device):
        _ffi.check(_ffi.lib().wf_eye_traces(_dev.ptr(rows), B, n_tr * span + 1 if n_tr else 0, T, span, _dev.ptr(out),
                                            _dev.stream_ptr(rows.device)))
    t = (time[: span + 1] - time[0]) + t_offset   # time is uniform: every trace shares this axis
    return t, (out[0] if was_1d else out)
