"""Headless equivalent of the reference's examples/irig_comparison.py:40-134: modulate the same bits as
PCM/FM, SOQPSK-TG and ARTM multi-h CPM (sps 20, the pulses of :50-72) and compare their power spectral
densities (Welch, NFFT 1024, amplitude sqrt(bits per symbol), frequency axis in units of the BIT rate, :127-134).
The reference draws two figures (pulses, PSD); figure output is unavailable on this box, so this script prints
the numbers behind them -- per waveform the pulse length, the 99 % and -25 dBc occupied bandwidths read off the
PSD, and optionally writes the curves (--out file.npz: freqs_<name>, psd_db_<name>, pulse_<name>, q_<name>).

    python examples/irig_comparison.py [--bits 65536] [--batch 8] [--out psd.npz]

All device work goes through the library: wf_prbs_encode, wf_cpm_modulate, wf_welch_psd."""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from waveforms_b200.cpm.modulate import cpm_modulate  # noqa: E402
from waveforms_b200.cpm.multih import MULTIH_IRIG_DENOM, MULTIH_IRIG_NUMER, freq_pulse_multih_irig  # noqa: E402
from waveforms_b200.cpm.pcmfm import PCMFM_DENOM, PCMFM_NUMER, freq_pulse_pcmfm  # noqa: E402
from waveforms_b200.cpm.soqpsk import SOQPSK_DENOM, SOQPSK_NUMER, freq_pulse_soqpsk_tg  # noqa: E402
from waveforms_b200.cpm.trellis.encoder import TrellisEncoder  # noqa: E402
from waveforms_b200.cpm.trellis.model import SimpleTrellis2, SimpleTrellis4, SOQPSKTrellis4x2DiffEncoded  # noqa: E402
from waveforms_b200.glfsr import PNSequence  # noqa: E402
from waveforms_b200.viz import plot_power_spectral_density  # noqa: E402


def occupied_bandwidth(freqs: np.ndarray, pxx: np.ndarray, fraction: float = 0.99) -> float:
    """Width of the smallest symmetric band around DC that holds ``fraction`` of the power."""
    order = np.argsort(np.abs(freqs))
    csum = np.cumsum(pxx[order])
    k = int(np.searchsorted(csum, fraction * csum[-1]))
    return 2.0 * float(np.abs(freqs[order][min(k, order.size - 1)]))


def width_at(freqs: np.ndarray, psd_db: np.ndarray, level_db: float) -> float:
    """Width of the band outside which the PSD stays below ``level_db`` (relative to its peak)."""
    above = np.flatnonzero(psd_db - psd_db.max() >= level_db)
    return float(freqs[above[-1]] - freqs[above[0]]) if above.size else 0.0


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--bits", type=int, default=65536, help="bits per frame")
    ap.add_argument("--batch", type=int, default=8, help="frames averaged into each PSD")
    ap.add_argument("--sps", type=int, default=20)
    ap.add_argument("--nfft", type=int, default=1024)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    sps = a.sps
    waveforms = [  # irig_comparison.py:50-72
        ("PCM-FM", TrellisEncoder(SimpleTrellis2), [PCMFM_NUMER / PCMFM_DENOM], freq_pulse_pcmfm(sps=sps, order=6), 1),
        ("SOQPSK-TG", TrellisEncoder(SOQPSKTrellis4x2DiffEncoded), [SOQPSK_NUMER / SOQPSK_DENOM], freq_pulse_soqpsk_tg(sps=sps), 1),
        ("Multi-h CPM", TrellisEncoder(SimpleTrellis4), MULTIH_IRIG_NUMER / MULTIH_IRIG_DENOM, freq_pulse_multih_irig(sps=sps), 2),
    ]
    bits = PNSequence(15).frames(first_frame=0, batch=a.batch, nbits=a.bits, device=dev)
    report, curves = {}, {}
    for name, mapper, h, pulse, bps in waveforms:
        symbols = mapper(bits)
        _t, signal = cpm_modulate(symbols, h, pulse, sps)
        freqs, pxx = plot_power_spectral_density(signal, sps, bps, a.nfft)
        pxx = pxx.mean(dim=0).double().cpu().numpy()          # average the frames' periodograms
        psd_db = 10.0 * np.log10(np.maximum(pxx, 1e-30))
        report[name] = {"pulse_symbol_times": pulse.size / sps, "bits_per_symbol": bps,
                        "bw99_bit_rates": occupied_bandwidth(freqs, pxx, 0.99),
                        "bw_minus25dBc_bit_rates": width_at(freqs, psd_db, -25.0),
                        "peak_db": float(psd_db.max())}
        key = name.lower().replace("-", "").replace(" ", "_")
        curves.update({f"freqs_{key}": freqs, f"psd_db_{key}": psd_db, f"pulse_{key}": pulse, f"q_{key}": np.cumsum(pulse) / sps})
    if a.out:
        np.savez_compressed(a.out, **curves)
    print(json.dumps(report, indent=1))


if __name__ == "__main__":
    main()
