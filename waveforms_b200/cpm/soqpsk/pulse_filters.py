"""SOQPSK frequency pulses (host, float64, set-up time) -- mirrors
waveforms/cpm/soqpsk/pulse_filters.py:5-116.  IRIG 106 / Perrins parameterisation:
g(t) ~ cos(pi rho B t / 2) / (1 - (rho B t)^2) * sinc(B t / 2) under a raised-cosine
window that is 1 for |t| < 2 T1 and falls to 0 at |t| = 2 (T1 + T2); scaled so the
taps sum to sps/2."""
from __future__ import annotations

import numpy as np

SOQPSK_NUMER = 1
SOQPSK_DENOM = 4


def freq_pulse_soqpsk(t1: float = 1.5, t2: float = 0.5, rho: float = 0.7, b: float = 1.25, sps: int = 8) -> np.ndarray:
    half_span = 2.0 * (t1 + t2)
    n_taps = int(2 * half_span * sps) + 1
    t = np.linspace(-half_span, half_span, num=n_taps, dtype=np.float64)
    core = np.cos(np.pi * rho * b * t / 2) / (1 - np.power(rho * b * t, 2)) * np.sinc(b * t / 2)
    window = np.ones_like(t)
    if t2 > 0:
        at = np.abs(t)
        taper = (at >= 2 * t1) & (at <= half_span)
        window[taper] = 0.5 * (1 + np.cos(np.pi * (t[taper] / 2 - t1) / t2))
        window[at > half_span] = 0
    shaped = core * window
    return core * (sps / (2 * np.sum(shaped))) * window


def freq_pulse_soqpsk_tg(sps: int = 8) -> np.ndarray:
    """SOQPSK-TG: B=1.25, T1=1.5, T2=0.5, rho=0.7 (8*sps+1 taps)."""
    return freq_pulse_soqpsk(t1=1.5, t2=0.5, rho=0.7, b=1.25, sps=sps)


def freq_pulse_soqpsk_a(sps: int = 8) -> np.ndarray:
    """SOQPSK-A: B=1.35, T1=1.4, T2=0.6, rho=1.0."""
    return freq_pulse_soqpsk(t1=1.4, t2=0.6, rho=1.0, b=1.35, sps=sps)


def freq_pulse_soqpsk_b(sps: int = 8) -> np.ndarray:
    """SOQPSK-B: B=1.45, T1=2.8, T2=1.2, rho=0.5 (16*sps+1 taps)."""
    return freq_pulse_soqpsk(t1=2.8, t2=1.2, rho=0.5, b=1.45, sps=sps)


def freq_pulse_soqpsk_mil(sps: int = 8) -> np.ndarray:
    """SOQPSK-MIL: one-symbol rectangle, sps+1 taps, g[0]=0 and the rest 1/2."""
    g = np.full(sps + 1, 0.5, dtype=np.float64)
    g[0] = 0.0
    return g
