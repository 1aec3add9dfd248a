"""Matched-filter banks for SOQPSK detection -- library home of the inline code at
examples/soqpsk_detection.py:57-63 (pseudo-symbols), :134-156 (pulse-truncation bank),
:158-173 (PAM bank) and the driver's symbol-instant pick (:189-196).

Filter design is host float64 (set-up); filtering is the decimated wf_mf_bank kernel,
which produces only the columns the detector reads: one 3-vector (alpha = -2, 0, +2)
per detector call."""
from __future__ import annotations

import numpy as np
import torch

from . import _dev, _ffi
from .cpm.pamapprox import rho_pulses

SQRT_HALF = np.sqrt(2) / 2
# pseudo-symbols of the two-pulse PAM approximation, rows k = 0, 1, columns alpha = -2, 0, +2
PSEUDO_SYMBOLS = np.array([[-1j, 1, 1j], [SQRT_HALF * (1 - 1j), SQRT_HALF, SQRT_HALF * (1 + 1j)]], dtype=np.complex128)
SOQPSK_ALPHABET = (-2, 0, 2)


def pt_taps(pulse_filter, mod_index: float, sps: int) -> np.ndarray:
    """Pulse-truncation filters (soqpsk_detection.py:135-156): exp(-2j pi h alpha q(t)) over
    the middle symbol of the phase pulse q, sps+1 taps, one row per alpha."""
    g = np.asarray(pulse_filter, dtype=np.float64)
    L = int(g.size / sps)
    q = np.cumsum(g) / sps
    mid = q[int((L - 1) * sps / 2):int((L + 1) * sps / 2) + 1]
    return np.stack([np.exp(-2j * np.pi * mod_index * a * mid) for a in SOQPSK_ALPHABET])


def pam_taps(pulse_filter, mod_index: float, sps: int):
    """(taps real[2, d_max], combine complex[3, 2]) of the PAM bank (soqpsk_detection.py:158-173):
    rho_0 / rho_1 zero-padded AT THE END to a common length, z_alpha = sum_k y_k conj(P[k, alpha])."""
    rho = rho_pulses(pulse_filter, mod_index, sps, k_max=2)
    d_max = max(r.size for r in rho)
    taps = np.stack([np.concatenate((r, np.zeros(d_max - r.size))) for r in rho])
    return taps, np.conj(PSEUDO_SYMBOLS).T.copy()


def symbol_instants(n_samples: int, sps: int, timing_offset: int = 0, length: int = 2):
    """(first_instant, n_calls) of the driver loop soqpsk_detection.py:189-192:
    n in [0, n_samples - length*sps) with (n + timing_offset) % sps == 0."""
    first = (-timing_offset) % sps
    limit = n_samples - length * sps
    n_calls = 0 if limit <= first else (limit - 1 - first) // sps + 1
    return first, n_calls


def filter_bank(rx: torch.Tensor, taps, combine=None, *, sps: int, first_instant: int, n_calls: int,
                center: int | None = None, out: torch.Tensor | None = None, n_samples: int | None = None) -> torch.Tensor:
    """Decimated bank on complex[B, ld] (or [T]) -> complex[B, n_calls, n_out].  ``n_samples`` is the
    valid signal length T when the rows of ``rx`` are padded."""
    rx = _dev.require_cuda(rx, "rx")
    if rx.dtype not in _dev.COMPLEX:
        raise TypeError("rx must be complex64 or complex128")
    rows, was_1d = _dev.as_batch(rx)
    B, ld = rows.shape
    T = ld if n_samples is None else int(n_samples)
    taps = np.asarray(taps)
    n_filt, n_taps = taps.shape
    tr = _dev.host_f64(taps.real.ravel())
    ti = _dev.host_f64(taps.imag.ravel()) if np.iscomplexobj(taps) else None
    cr = ci = None
    n_out = n_filt
    if combine is not None:
        combine = np.asarray(combine, dtype=np.complex128)
        n_out = combine.shape[0]
        cr, ci = _dev.host_f64(combine.real.ravel()), _dev.host_f64(combine.imag.ravel())
    if center is None:
        center = (n_taps - 1) // 2
    if out is None:
        ld_out = n_calls + (n_calls & 1)
        out = torch.empty((B, ld_out, n_out), dtype=rows.dtype, device=rows.device)
    else:
        ld_out = out.shape[1]
    with torch.cuda.device(rows.device):
        _ffi.check(_ffi.lib().wf_mf_bank(_dev.ptr(rows), _dev.COMPLEX[rows.dtype], B, T, ld, _dev.f64_ptr(tr),
                                         None if ti is None else _dev.f64_ptr(ti), n_filt, n_taps, int(center),
                                         None if cr is None else _dev.f64_ptr(cr),
                                         None if ci is None else _dev.f64_ptr(ci), n_out, int(sps), int(first_instant),
                                         int(n_calls), _dev.ptr(out), ld_out, _dev.stream_ptr(rows.device)))
    z = out[:, :n_calls]
    return z[0] if was_1d else z


def pt_bank(rx, pulse_filter, mod_index: float, sps: int, *, timing_offset: int = 0, length: int = 2, **kw):
    rows = rx if rx.dim() == 2 else rx.unsqueeze(0)
    T = kw.get("n_samples") or rows.shape[1]
    first, n_calls = symbol_instants(T, sps, timing_offset, length)
    return filter_bank(rx, pt_taps(pulse_filter, mod_index, sps), sps=sps, first_instant=first, n_calls=n_calls, **kw)


def pam_bank(rx, pulse_filter, mod_index: float, sps: int, *, timing_offset: int = 0, length: int = 2, **kw):
    rows = rx if rx.dim() == 2 else rx.unsqueeze(0)
    T = kw.get("n_samples") or rows.shape[1]
    first, n_calls = symbol_instants(T, sps, timing_offset, length)
    taps, comb = pam_taps(pulse_filter, mod_index, sps)
    return filter_bank(rx, taps, comb, sps=sps, first_instant=first, n_calls=n_calls, **kw)
