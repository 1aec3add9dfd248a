"""Headless equivalent of the reference's examples/soqpsk_detection.py:35-216, written against the module API the
reference exposes (same names, a batch axis added): PN15 bits -> SOQPSKTrellis4x2DiffEncoded -> cpm_modulate
(SOQPSK-MIL and -TG, sps 10) -> complex AWGN at sigma = sqrt(2)/2 (Eb/N0 = 10 dB) -> PT and PAM matched-filter
banks at the script's timing offsets (:183-187) -> SOQPSKTrellisDetector -> error count (:200-216).

The reference runs ONE frame and draws eye diagrams / constellations; here --frames frames run at once with
independent Philox noise, the error counts are printed as JSON, and the arrays behind the figures (eye traces of
the noisy baseband, constellation of the detector inputs at the symbol instants) can be written with --out
(figure output is unavailable on this box).

    python examples/soqpsk_detection.py [--frames 64] [--sigma 0.7071] [--out arrays.npz]
"""
from __future__ import annotations

import argparse
import cmath
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from waveforms_b200 import bert, mfbank  # noqa: E402
from waveforms_b200.cpm.modulate import cpm_modulate  # noqa: E402
from waveforms_b200.cpm.soqpsk import freq_pulse_soqpsk_mil, freq_pulse_soqpsk_tg  # noqa: E402
from waveforms_b200.cpm.trellis.encoder import TrellisEncoder  # noqa: E402
from waveforms_b200.cpm.trellis.model import SOQPSKTrellis4x2DiffEncoded  # noqa: E402
from waveforms_b200.glfsr import PNSequence  # noqa: E402
from waveforms_b200.noise import PhiloxRNG, generate_complex_awgn  # noqa: E402
from waveforms_b200.viterbi import SOQPSKTrellisDetector  # noqa: E402
from waveforms_b200.viz import plot_eye_diagram  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--bits", type=int, default=32768, help="bits per frame (the example: 4096 bytes of PN15)")
    ap.add_argument("--sps", type=int, default=10)
    ap.add_argument("--sigma", type=float, default=math.sqrt(2) / 2)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    sps, B, h = a.sps, a.frames, 0.25
    ebn0_db = 10 * math.log10(sps / (2 * a.sigma**2))                                  # soqpsk_detection.py:132
    bits = PNSequence(15).frames(first_frame=0, batch=B, nbits=a.bits, device=dev)     # :28-30,45
    symbols = TrellisEncoder(SOQPSKTrellis4x2DiffEncoded)(bits)                         # :48-49
    sigma = torch.full((B,), a.sigma, dtype=torch.float32, device=dev)
    report, arrays = {"ebn0_db": ebn0_db, "frames": B, "bits_per_frame": a.bits}, {}
    # (label, pulse, {detector: timing offset}) as hard-coded at :183-187
    for k, (label, pulse_fn, offsets) in enumerate((("MIL", freq_pulse_soqpsk_mil, {"PT": -1, "PAM": -3}),
                                                    ("TG", freq_pulse_soqpsk_tg, {"PT": -1, "PAM": 0}))):
        pulse = pulse_fn(sps)
        _t, tx = cpm_modulate(symbols, h, pulse, sps)                                   # :78-83
        noise = generate_complex_awgn(sigma, tx.shape[1], PhiloxRNG(a.seed + k, first_frame=0), batch=B)
        rx = tx * cmath.exp(-1j * math.pi / 4) + noise                                  # :84-90
        for det_name, off in offsets.items():
            bank = mfbank.pt_bank if det_name == "PT" else mfbank.pam_bank             # :134-173, instants :189-196
            z = bank(rx, pulse, h, sps, timing_offset=off)
            bits_hat, _syms_hat = SOQPSKTrellisDetector(2, differantial_encoding=True).detect(z)   # :177-198
            errors, _sym_errors, compared = bert.count_errors(bits_hat, bits, skip=2)  # :200-209
            report[f"{label}_{det_name}"] = {"timing_offset": off, "bit_errors": errors, "compared": compared,
                                             "ber": errors / max(compared, 1)}
            if a.out:
                arrays[f"constellation_{label}_{det_name}"] = z[0, :1024, 1].cpu().numpy()  # alpha = 0 branch, frame 0
        if a.out:
            _te, traces = plot_eye_diagram(_t, rx[:1].contiguous(), sps, modulo=4)
            arrays[f"eye_{label}"] = traces[0, :256].cpu().numpy()
    if a.out:
        np.savez_compressed(a.out, **arrays)
    print(json.dumps(report, indent=1))


if __name__ == "__main__":
    main()
