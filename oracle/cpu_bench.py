"""CPU baseline driver (TEST/BENCH INFRASTRUCTURE ONLY): the oracle's restatement of the
reference chain (examples/soqpsk_detection.py:45-216) run over many frames on all host
cores.  Used only by bench.py's ``cpu_baseline`` leg and ``--impl reference`` arm.

kind = "port": the reference itself is pure Python + numpy and cannot travel to the GPU
box, so the timed code is the oracle port -- same float64 arithmetic per frame
(np.convolve modulator and banks), with the detector vectorised across the frames of a
chunk (decision-identical to the reference's per-call detector; asserted in tests).
That makes this baseline FASTER than the unmodified reference (about 1.8e4 bits/s/core
measured in the survey), i.e. conservative for any speed-up quoted against it."""
from __future__ import annotations

import multiprocessing as mp
import os
import time

import numpy as np

from . import cpm_link as o

_PULSE = {}


def _chunk(args):
    first_frame, n_frames, n_bits, sps, ebn0_db, detector = args
    key = sps
    if key not in _PULSE:
        _PULSE[key] = o.pulse_soqpsk_tg(sps)
    pulse = _PULSE[key]
    sigma = float(o.sigma_from_ebn0_db(ebn0_db, sps))
    cols = o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"]
    zs, bits_all, syms_all = [], [], []
    for f in range(first_frame, first_frame + n_frames):
        bits = o.pn_frame_bits(15, f, n_bits)
        syms = o.fsm_encode(cols, bits)
        s = o.cpm_modulate_closed_form(syms, 0.25, pulse, sps)
        rng = np.random.Generator(np.random.PCG64(seed=1000 * int(round(ebn0_db)) + f))
        rx = s * np.exp(-1j * np.pi / 4) + o.complex_awgn(sigma, s.size, rng)
        mf = o.pt_bank(rx, pulse, 0.25, sps) if detector == "PT" else o.pam_bank(rx, pulse, 0.25, sps)
        inst = o.symbol_instants(rx.size, sps, 0, 2)
        zs.append(mf[:, inst].T)
        bits_all.append(bits)
        syms_all.append(syms)
    db, ds = o.detect_soqpsk_fast(np.stack(zs), True)
    be = n = 0
    for k in range(n_frames):
        e, _se, c = o.count_errors(db[k], ds[k], bits_all[k], syms_all[k], 2)
        be += e
        n += c
    return be, n


def run(n_frames: int, n_bits: int = 4096, sps: int = 8, ebn0_db: float = 6.0, detector: str = "PAM",
        cores: int | None = None, first_frame: int = 0, frames_per_task: int = 16):
    """Simulate n_frames frames on ``cores`` processes.  Returns dict(bits, seconds, bits_per_s,
    bit_errors, compared, cores)."""
    cores = cores or os.cpu_count() or 1
    tasks = [(first_frame + f0, min(frames_per_task, n_frames - f0), n_bits, sps, ebn0_db, detector)
             for f0 in range(0, n_frames, frames_per_task)]
    t0 = time.perf_counter()
    if cores == 1:
        res = [_chunk(t) for t in tasks]
    else:
        with mp.get_context("fork").Pool(cores) as pool:
            res = pool.map(_chunk, tasks, chunksize=1)
    dt = time.perf_counter() - t0
    be = sum(r[0] for r in res)
    n = sum(r[1] for r in res)
    return dict(bits=n_frames * n_bits, seconds=dt, bits_per_s=n_frames * n_bits / dt, bit_errors=be, compared=n,
                cores=cores)


def run_for(seconds: float, **kw):
    """Size the sample so the run takes about ``seconds``: a short pilot, then the real run."""
    cores = kw.get("cores") or os.cpu_count() or 1
    per = kw.get("frames_per_task", 16)
    pilot = run(cores * per, **kw)
    n = int(max(cores * per, pilot["bits_per_s"] * seconds / kw.get("n_bits", 4096)))
    n = (n // (cores * per)) * cores * per or cores * per
    out = run(n, **{**kw, "first_frame": cores * per})
    out["n_frames"] = n
    return out
