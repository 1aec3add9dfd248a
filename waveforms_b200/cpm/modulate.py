"""CPM modulators on the GPU -- mirrors waveforms/cpm/modulate.py:12-101.

``cpm_modulate(symbols, mod_index, pulse_filter, sps)`` keeps the reference signature;
``symbols`` is an int8 CUDA tensor, [S] for one frame or [B, S] for a batch of
independent frames.  The result is complex64 by default (``dtype=torch.complex128``
runs the float64 instantiation used by the parity tests).  All per-sample work runs
in wf_cpm_modulate (closed-form phase, no sequential scan over samples)."""
from __future__ import annotations

import math

import numpy as np
import torch

from .. import _dev, _ffi

INITIAL_PHASE = math.pi / 4  # modulate.py:100


def _as_real_f64(t, name):
    t = _dev.require_cuda(t, name)
    return t.to(torch.float64).contiguous()


def phase_modulate(phase: torch.Tensor, sensitivity: float, *, dtype=torch.complex128) -> torch.Tensor:
    """exp(j * sensitivity * phase) (modulate.py:12-25)."""
    ph = _as_real_f64(phase, "phase")
    out = torch.empty(ph.shape, dtype=dtype, device=ph.device)
    with torch.cuda.device(ph.device):
        _ffi.check(_ffi.lib().wf_phase_modulate(_dev.ptr(ph), ph.numel(), float(sensitivity), _dev.COMPLEX[dtype],
                                                _dev.ptr(out), _dev.stream_ptr(ph.device)))
    return out


def frequency_modulate(freq_pulses: torch.Tensor, sps: int, initial_phase: float = 0, *,
                       dtype=torch.complex128) -> torch.Tensor:
    """Accumulate-with-modulo FM of pulse-shaped symbols (modulate.py:28-54); rows are frames."""
    f = _as_real_f64(freq_pulses, "freq_pulses")
    rows, was_1d = _dev.as_batch(f)
    B, T = rows.shape
    out = torch.empty((B, T), dtype=dtype, device=f.device)
    with torch.cuda.device(f.device):
        _ffi.check(_ffi.lib().wf_frequency_modulate(_dev.ptr(rows), B, T, T, int(sps), float(initial_phase),
                                                    _dev.COMPLEX[dtype], _dev.ptr(out), T,
                                                    _dev.stream_ptr(f.device)))
    return out[0] if was_1d else out


def modulate_frames(symbols: torch.Tensor, mod_index, pulse_filter, sps: int = 8, *, dtype=torch.complex64,
                    initial_phase: float = INITIAL_PHASE, derotate: float = 0.0, sigma: torch.Tensor | None = None,
                    point: torch.Tensor | None = None, seed: int = 0, first_frame: int = 0,
                    out: torch.Tensor | None = None, frame_ids: torch.Tensor | None = None) -> torch.Tensor:
    """Batched core: int8[B, S] -> complex[B, T], T = (S+1)*sps, optionally with AWGN of
    per-frame ``sigma`` (float32[B]) fused in.  ``out`` may be a preallocated [B, ld>=T] buffer;
    ``frame_ids`` (int64[B], optional) keys every row's noise stream by its own global frame id."""
    symbols = _dev.require_cuda(symbols, "symbols", torch.int8)
    B, S = symbols.shape
    T = (S + 1) * sps
    h = _dev.host_f64(mod_index)
    g = _dev.host_f64(pulse_filter)
    if out is None:
        ld = T + (T & 1)
        out = torch.empty((B, ld), dtype=dtype, device=symbols.device)
    else:
        out = _dev.require_cuda(out, "out")
        dtype, ld = out.dtype, out.shape[1]
    if sigma is not None:
        sigma = _dev.require_cuda(sigma, "sigma", torch.float32)
    if point is not None:
        point = _dev.require_cuda(point, "point", torch.int32)
    if frame_ids is not None:
        frame_ids = _dev.require_cuda(frame_ids, "frame_ids", torch.int64)
    with torch.cuda.device(symbols.device):
        _ffi.check(_ffi.lib().wf_cpm_modulate(_dev.ptr(symbols), B, S, _dev.f64_ptr(h), h.size, _dev.f64_ptr(g), g.size,
                                              int(sps), float(initial_phase), float(derotate), _dev.COMPLEX[dtype],
                                              _dev.ptr(out), ld, _dev.ptr(sigma), _dev.ptr(point),
                                              int(seed) & (2**64 - 1), int(first_frame), _dev.ptr(frame_ids),
                                              _dev.stream_ptr(symbols.device)))
    return out[:, :T]


def cpm_modulate(symbols: torch.Tensor, mod_index, pulse_filter, sps: int = 8, *, dtype=torch.complex64):
    """Continuous-phase modulate ``symbols`` (modulate.py:57-101).

    Returns ``(normalized_time, modulated_signal)``: time in symbol periods, float64[T]
    (generated on the fly, one copy for the whole batch), and the unit-amplitude baseband
    signal, [T] or [B, T] following the shape of ``symbols``."""
    if isinstance(mod_index, (int, float)):
        mod_index = [float(mod_index)]
    symbols = _dev.require_cuda(symbols, "symbols", torch.int8)
    frames, was_1d = _dev.as_batch(symbols)
    sig = modulate_frames(frames, mod_index, pulse_filter, sps, dtype=dtype)
    n_pts = sig.shape[1]
    # np.linspace(0, S+1, n_pts, endpoint=False) == arange(n_pts) * ((S+1)/n_pts)  (modulate.py:83-88)
    step = (frames.shape[1] + 1) / n_pts
    normalized_time = torch.arange(n_pts, dtype=torch.float64, device=sig.device) * step
    return normalized_time, (sig[0] if was_1d else sig)
