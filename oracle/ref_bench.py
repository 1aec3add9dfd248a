"""Time the UNMODIFIED reference on this machine's host cores.

The reference package is imported from, in this order: $WAVEFORMS_REFERENCE, <repo>/baseline/_ref (the
git-ignored `pip install --no-index --no-deps --target baseline/_ref /root/reference` staged by
__graft_entry__.build() in the build container, which travels to the GPU box), /root/reference.  Every
numerical step is one of the reference's own functions (PNSequence, TrellisEncoder, cpm_modulate,
generate_complex_awgn, rho_pulses, SOQPSKTrellisDetector.iteration); the glue between them is the headless
replay of examples/soqpsk_detection.py (the script itself imports matplotlib and cannot run): bank :134-173,
driver loop :177-198, tally :200-209 -- WITHOUT the margin bookkeeping oracle/make_golden.py adds, and with
only the bank of the detector being timed, so the number is the reference at its best.

    python oracle/ref_bench.py [--frames-single 2] [--frames-pool-per-core 2] [--detector PAM] [--ebn0 6]

prints one JSON object: bits/s single-process (upstream is single-threaded) and with
multiprocessing.Pool(all cores) over independent frames, core count and CPU model stated.
Test / measurement infrastructure only (bench.py's cpu_baseline_reference leg); never on the product path."""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
N_BITS, SPS = 4096, 8


def locate():
    for cand in (os.environ.get("WAVEFORMS_REFERENCE"), os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if cand and os.path.isdir(os.path.join(cand, "waveforms")):
            return cand
    return None


def _imports():
    ref = locate()
    if ref is None:
        raise RuntimeError("reference package not found (baseline/_ref not staged and /root/reference absent)")
    if ref not in sys.path:
        sys.path.insert(0, ref)
    import waveforms  # noqa: F401
    assert os.path.abspath(waveforms.__file__).startswith(os.path.abspath(ref)), waveforms.__file__
    from waveforms.cpm.modulate import cpm_modulate
    from waveforms.cpm.pamapprox import rho_pulses
    from waveforms.cpm.soqpsk import freq_pulse_soqpsk_tg
    from waveforms.cpm.trellis import model as tm
    from waveforms.cpm.trellis.encoder import TrellisEncoder
    from waveforms.glfsr import PNSequence
    from waveforms.noise import generate_complex_awgn
    from waveforms.viterbi.algorithm import SOQPSKTrellisDetector
    return locals()


_R = None
_SEQ = None


def _frame(task):
    """One frame through the reference chain; returns (bit errors, compared bits)."""
    global _R, _SEQ
    f, ebn0, detector = task
    if _R is None:
        _R = _imports()
        _SEQ = np.array(_R["PNSequence"](15).generate_sequence(), dtype=np.uint8)  # soqpsk_detection.py:45-47
    R = _R
    bits = _SEQ[(f * N_BITS + np.arange(N_BITS)) % _SEQ.size]
    pulse, h = R["freq_pulse_soqpsk_tg"](SPS), 0.25
    symbols = R["TrellisEncoder"](R["tm"].SOQPSKTrellis4x2DiffEncoded)(bits)                  # :48-49
    _t, sig = R["cpm_modulate"](symbols=symbols, mod_index=h, pulse_filter=pulse, sps=SPS)   # :78-83
    sigma = float(np.sqrt(SPS / (2 * 10 ** (ebn0 / 10))))
    noise = R["generate_complex_awgn"](sigma, sig.size, np.random.Generator(np.random.PCG64(1000 * int(ebn0) + f)))
    rx = sig * np.exp(-1j * np.pi / 4) + noise                                               # :85-90
    mf = np.zeros((3, rx.size), dtype=np.complex128)
    if detector == "PT":                                                                     # :134-156
        L = int(pulse.size / SPS)
        q = np.cumsum(pulse) / SPS
        seg = q[int((L - 1) * SPS / 2): int((L + 1) * SPS / 2) + 1]
        for r, a in enumerate((-2, 0, 2)):
            mf[r] = np.convolve(rx, np.exp(-2j * np.pi * h * a * seg), mode="same")
    else:                                                                                    # :158-173
        pseudo = np.array([[-1j, 1, 1j], [np.sqrt(2) / 2 * (1 - 1j), np.sqrt(2) / 2, np.sqrt(2) / 2 * (1 + 1j)]])
        rho = R["rho_pulses"](pulse, h, SPS, k_max=2)
        d_max = max(r.size for r in rho)
        for s in range(3):
            for k in range(2):
                rk = np.concatenate((rho[k], np.zeros(d_max - rho[k].size)))
                mf[s] += np.convolve(rx, rk, mode="same") * np.conj(pseudo[k, s])
    det = R["SOQPSKTrellisDetector"](length=2)                                               # :177-198
    ob = []
    for n in range(rx.size - det.length * SPS):
        if n % SPS:
            continue
        rb, _rs = det.iteration(mf[:, n])
        ob.append(rb[0])
    b = np.array(ob[2:], dtype=np.uint8)                                                     # :200-209
    n = min(bits.size, b.size)
    return int(np.count_nonzero(b[:n] != bits[:n])), n


def run(n_frames, procs, ebn0=6.0, detector="PAM", first_frame=0):
    tasks = [(first_frame + f, ebn0, detector) for f in range(n_frames)]
    if procs <= 1:
        _frame(tasks[0])  # warm-up: imports, PN sequence
        t0 = time.perf_counter()
        res = [_frame(t) for t in tasks]
        dt = time.perf_counter() - t0
    else:
        with mp.get_context("fork").Pool(procs) as pool:
            pool.map(_frame, [(0, ebn0, detector)] * procs, chunksize=1)  # warm every worker
            t0 = time.perf_counter()
            res = pool.map(_frame, tasks, chunksize=1)
            dt = time.perf_counter() - t0
    return {"frames": n_frames, "procs": procs, "seconds": dt, "bits_per_s": n_frames * N_BITS / dt,
            "bit_errors": int(sum(r[0] for r in res)), "compared": int(sum(r[1] for r in res))}


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def measure(frames_single=2, frames_pool_per_core=2, ebn0=6.0, detector="PAM"):
    cores = os.cpu_count() or 1
    single = run(frames_single, 1, ebn0, detector)
    pool = run(frames_pool_per_core * cores, cores, ebn0, detector, first_frame=1000)
    return {"kind": "reference", "where": locate(), "cores": cores, "cpu_model": cpu_model(), "detector": detector,
            "ebn0_db": ebn0, "n_bits": N_BITS, "sps": SPS, "single_process": single, "pool_all_cores": pool,
            "unit": "bits/s"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames-single", type=int, default=2)
    ap.add_argument("--frames-pool-per-core", type=int, default=2)
    ap.add_argument("--detector", default="PAM")
    ap.add_argument("--ebn0", type=float, default=6.0)
    a = ap.parse_args()
    print(json.dumps(measure(a.frames_single, a.frames_pool_per_core, a.ebn0, a.detector)))
