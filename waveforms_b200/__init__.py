"""waveforms_b200 -- B200-native (sm_100a) implementation of the Monte-Carlo link
simulation hot path of mcdiarmid/waveforms, behind the reference's Python API:

    waveforms_b200.glfsr     GLFSR, PNSequence                (waveforms/glfsr)
    waveforms_b200.cpm       trellis encoder, pulses, mappers, cpm_modulate, rho_pulses
    waveforms_b200.noise     generate_complex_awgn            (waveforms/noise.py)
    waveforms_b200.mfbank    PT / PAM matched-filter banks    (examples/soqpsk_detection.py:134-173)
    waveforms_b200.viterbi   SOQPSKTrellisDetector            (waveforms/viterbi)
    waveforms_b200.bert      count_errors                     (examples/soqpsk_detection.py:200-209)
    waveforms_b200.linksim   LinkSimulator: the batched chain + multi-GPU sweep
    waveforms_b200.viz       PSD / eye / constellation / phase-tree ARRAYS (waveforms/viz, figures unavailable)
    waveforms_b200.filters   kaiser_fir_lpf + batched FIR + sampling-phase estimate (waveforms/filters)

Per-frame work runs in hand-written CUDA kernels (libwfb200.so, C ABI in
include/wfb200.h) on torch CUDA tensors; there is no CPU fallback.
``install_as("waveforms")`` aliases the package under the reference's import name."""
from __future__ import annotations

import importlib
import sys

__version__ = "0.1.0"

_SUBMODULES = (
    "glfsr", "glfsr.glfsr", "glfsr.pn", "cpm", "cpm.helpers", "cpm.modulate", "cpm.pamapprox", "cpm.trellis",
    "cpm.trellis.model", "cpm.trellis.encoder", "cpm.soqpsk", "cpm.soqpsk.precoder", "cpm.soqpsk.pulse_filters",
    "cpm.pcmfm", "cpm.pcmfm.precoder", "cpm.pcmfm.pulse_filters", "cpm.multih", "cpm.multih.precoder",
    "cpm.multih.pulse_filters", "noise", "viterbi", "viterbi.algorithm", "bert", "mfbank", "linksim", "viz", "viz.psd", "viz.eye",
    "viz.constellation", "viz.tree", "filters", "filters.lpf",
)


def install_as(name: str = "waveforms") -> None:
    """Register this package and its submodules in sys.modules under ``name`` so that
    ``from waveforms.cpm.modulate import cpm_modulate`` resolves to the CUDA implementation."""
    pkg = sys.modules[__name__]
    sys.modules[name] = pkg
    for sub in _SUBMODULES:
        sys.modules[f"{name}.{sub}"] = importlib.import_module(f"{__name__}.{sub}")
