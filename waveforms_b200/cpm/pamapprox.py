"""PAM (Laurent / Perrins) decomposition pulses, host float64 set-up -- mirrors
waveforms/cpm/pamapprox.py:12-102.  rho_k feed the PAM matched-filter bank
(waveforms_b200.mfbank.pam_bank)."""
from __future__ import annotations

import numpy as np


def pam_unit_pulse(phase_pulse: np.ndarray, mod_index: float) -> np.ndarray:
    """u(t) on 2*len+1 samples with a leading zero (pamapprox.py:12-29)."""
    q = np.asarray(phase_pulse, dtype=np.float64)
    pih = mod_index * np.pi
    u = np.zeros(2 * q.size + 1, dtype=np.float64)
    u[1:q.size + 1] = np.sin(2 * pih * q) / np.sin(pih)
    u[q.size + 1:] = np.sin(pih - 2 * pih * q) / np.sin(pih)
    return u


def pam_unit_pulse2(phase_pulse: np.ndarray, mod_index: float) -> np.ndarray:
    """u(t) on 2*len-1 samples (pamapprox.py:32-49): rising half sin(2 pi h q)/sin(pi h),
    falling half sin(pi h - 2 pi h q)/sin(pi h) without repeating the peak."""
    q = np.asarray(phase_pulse, dtype=np.float64)
    pih = mod_index * np.pi
    u = np.zeros(2 * q.size - 1, dtype=np.float64)
    u[:q.size] = np.sin(2 * pih * q) / np.sin(pih)
    u[q.size:] = np.sin(pih - 2 * pih * q[1:]) / np.sin(pih)
    return u


def rho_pulses(pulse_filter: np.ndarray, mod_index: float, sps: int, k_max: int = 2) -> list[np.ndarray]:
    """rho_k(t) = C_k * prod over 2L shifted copies of u(t) (Perrins eq. 3.21), cropped to
    its support, for k < k_max (pamapprox.py:52-102).  Row k of the shift table is
    {x + c*[k>0] : x in 0..L-1, c in 0..1} symbol periods, C_k = k+1."""
    g = np.asarray(pulse_filter, dtype=np.float64)
    L = int(g.size / sps)
    u = pam_unit_pulse2(np.cumsum(g) / sps, mod_index)
    pulses = []
    for k in range(k_max):
        shifts = [x + (c if k > 0 else 0) for x in range(L) for c in (0, 1)]
        prod = np.ones(u.size + L * sps, dtype=np.float64)
        for s in shifts:
            prod *= np.concatenate((np.zeros(s * sps), u, np.zeros((L - s) * sps)))
        pulses.append(((k + 1.0) * prod)[shifts[-1] * sps:-L * sps])
    return pulses
