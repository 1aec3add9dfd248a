"""Generate tests/golden/*.npz by running the REFERENCE ITSELF (read-only at
/root/reference) -- run once in the build container, outputs committed.

    python oracle/make_golden.py [--out tests/golden] [--skip-example]

The reference has no library home for the matched-filter banks, the detector
driver loop and the error count: they are inline in
examples/soqpsk_detection.py (which imports matplotlib and cannot run here).
``replay_frame`` below is a headless replay of those lines calling the
reference's own functions in the script's order; line numbers are cited.
Nothing here is imported by the product or by tests (tests read the .npz).
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

REF = os.environ.get("WAVEFORMS_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

from waveforms.cpm.modulate import cpm_modulate  # noqa: E402
from waveforms.cpm.multih import MULTIH_IRIG_DENOM, MULTIH_IRIG_NUMER, MultiHSymbolMapper, freq_pulse_multih_irig  # noqa: E402
from waveforms.cpm.pamapprox import rho_pulses  # noqa: E402
from waveforms.cpm.pcmfm import PCMFM_DENOM, PCMFM_NUMER, PCMFMSymbolMapper, freq_pulse_pcmfm  # noqa: E402
from waveforms.cpm.soqpsk import (  # noqa: E402
    SOQPSKPrecoder,
    freq_pulse_soqpsk_a,
    freq_pulse_soqpsk_b,
    freq_pulse_soqpsk_mil,
    freq_pulse_soqpsk_tg,
)
from waveforms.cpm.trellis import model as tm  # noqa: E402
from waveforms.cpm.trellis.encoder import TrellisEncoder  # noqa: E402
from waveforms.glfsr import PNSequence  # noqa: E402
from waveforms.glfsr.pn import generate_mask  # noqa: E402
from waveforms.noise import generate_complex_awgn  # noqa: E402
from waveforms.viterbi.algorithm import SOQPSKTrellisDetector  # noqa: E402

PSEUDO = np.array(
    [[-1j, 1, 1j], [np.sqrt(2) / 2 * (1 - 1j), np.sqrt(2) / 2, np.sqrt(2) / 2 * (1 + 1j)]],
    dtype=np.complex128,
)  # soqpsk_detection.py:57-63


def banks(rx, pulse, h, sps):
    """soqpsk_detection.py:134-173, both banks, full rate."""
    L = int(pulse.size / sps)
    q = np.cumsum(pulse) / sps
    seg = q[int((L - 1) * sps / 2): int((L + 1) * sps / 2) + 1]
    pt = np.zeros((3, rx.size), dtype=np.complex128)
    for r, a in enumerate((-2, 0, 2)):
        pt[r] = np.convolve(rx, np.exp(-2j * np.pi * h * a * seg), mode="same")
    rho = rho_pulses(pulse, h, sps, k_max=2)
    d_max = max(r.size for r in rho)
    pam = np.zeros((3, rx.size), dtype=np.complex128)
    for s in range(3):
        for k in range(2):
            rk = np.concatenate((rho[k], np.zeros(d_max - rho[k].size)))
            pam[s] += np.convolve(rx, rk, mode="same") * np.conj(PSEUDO[k, s])
    return pt, pam, rho


def acs_margin(det, carried, i):
    """Float64 decision margins of the detector call that just ran (call counter was ``i``,
    ``carried`` = metrics[:, 0] before the call), recomputed from the reference object's own
    state arrays (algorithm.py:65-90): for every stage j and end state the |difference| of the
    competing candidate metrics, and the gap between the two smallest final metrics that
    np.argmin separates (:92).  Returns (smallest margin of the call, largest |candidate|): a
    float32 detector may legitimately differ from the float64 one only where the first is within
    rounding of zero relative to the second."""
    n = det.length
    prev = carried - carried.min()
    small, scale = np.inf, 0.0
    for j in range(n):
        branches = det.fsm.trellis.branches[(i + j - 1) % det.fsm.columns]
        for st in range(det.fsm.states):
            c = [prev[b.start] + det.bi_history[k, j] for k, b in enumerate(branches) if b.end == st]
            scale = max(scale, max(abs(x) for x in c))
            c.sort()
            if len(c) > 1:
                small = min(small, c[1] - c[0])
        prev = det.metrics[:, j]
    fin = np.sort(det.metrics[:, -1])
    return min(small, fin[1] - fin[0]), scale


def drive(mf, n_samples, sps, timing_offset, length=2, diff=True):
    """soqpsk_detection.py:177-198."""
    det = SOQPSKTrellisDetector(length=length, differantial_encoding=diff)
    ob, osy, inst, marg = [], [], [], []
    for n in range(n_samples - det.length * sps):
        if (n + timing_offset) % sps:
            continue
        carried, i = det.metrics[:, 0].copy(), det.i
        rb, rs = det.iteration(mf[:, n])
        ob.append(rb[0])
        osy.append(rs[0])
        inst.append(n)
        marg.append(acs_margin(det, carried, i))
    return np.array(ob), np.array(osy), np.array(inst), np.array(marg)


def tally(ob, osy, bits, symbols, length=2):
    """soqpsk_detection.py:200-209."""
    s = np.array(osy[length:], dtype=np.int8)
    b = np.array(ob[length:], dtype=np.uint8)
    n = min(symbols.size, s.size)
    se = len(np.where(s[:n] - symbols[:n])[0])
    be = len(np.where(b[:n] - bits[:n])[0])
    return be, se, n


def replay_frame(bits, pulse, sps, sigma, rng, offsets, h=0.25, diff=True, length=2):
    """soqpsk_detection.py:48-49,78-90 then banks/drive/tally for PT and PAM."""
    trellis = tm.SOQPSKTrellis4x2DiffEncoded if diff else tm.SOQPSKTrellis4x2
    symbols = TrellisEncoder(trellis)(bits)
    _t, sig = cpm_modulate(symbols=symbols, mod_index=h, pulse_filter=pulse, sps=sps)
    tx = sig.copy()
    noise = generate_complex_awgn(sigma, sig.size, rng)
    sig[:] *= np.exp(-1j * np.pi / 4)
    rx = sig + noise
    pt, pam, rho = banks(rx, pulse, h, sps)
    out = dict(symbols=symbols, tx=tx, noise=noise, rho0=rho[0], rho1=rho[1])
    for name, off in offsets.items():  # "PT", "PAM", or e.g. "PT@-1" for a second timing offset on the same bank
        mf = pt if name.startswith("PT") else pam
        ob, osy, inst, marg = drive(mf, rx.size, sps, off, length, diff)
        be, se, n = tally(ob, osy, bits, symbols, length)
        out[name] = dict(z=mf[:, inst].T.copy(), dec_bits=ob.astype(np.uint8),
                         dec_syms=osy.astype(np.int8), bit_errors=be, sym_errors=se, compared=n,
                         margin=marg[:, 0].copy(), scale=marg[:, 1].copy())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(os.path.dirname(__file__), "..", "tests", "golden"))
    ap.add_argument("--skip-example", action="store_true")
    ap.add_argument("--only-trellis", action="store_true", help="write only g9_trellis.npz")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    sv = lambda name, **kw: np.savez_compressed(os.path.join(a.out, name), **kw)  # noqa: E731

    # G9: the reference's trellis tables, branch order included (model.py:179-293)
    g9 = {}
    for name in ("SOQPSKTrellis8x1", "SOQPSKTrellis4x2", "SOQPSKTrellis4x2DiffEncoded", "SimpleTrellis2", "SimpleTrellis4"):
        t = getattr(tm, name)
        g9[name] = np.array([[(b.inp, b.out, b.start, b.end) for b in col] for col in t.branches], dtype=np.int64)
        g9[name + "_shape"] = np.array([t.columns, t.states, t.branches_per_column, t.input_cardinality,
                                        t.output_cardinality])
    sv("g9_trellis.npz", **g9)
    if a.only_trellis:
        return

    # G1: PN bits + masks
    g1 = {"masks": np.array([generate_mask(d) for d in range(2, 65)], dtype=np.uint64)}
    for d in (9, 15):
        g1[f"pn{d}"] = np.packbits(np.array(PNSequence(d).generate_sequence(), dtype=np.uint8))
    p23 = PNSequence(23)
    g1["pn23_first65536"] = np.packbits(np.array([p23.next_bit() for _ in range(65536)], dtype=np.uint8))
    for d in (2, 3, 5, 7, 12, 16, 31, 32, 33, 63, 64):
        p = PNSequence(d)
        g1[f"pn{d}_first256"] = np.packbits(np.array([p.next_bit() for _ in range(256)], dtype=np.uint8))
    sv("g1_pn.npz", **g1)

    # bit stream of the examples: PN15 packed to 4096 bytes (one pad bit)
    bit_array = np.unpackbits(np.packbits(PNSequence(15).generate_sequence()))

    # G2: encoders / mappers on the 32768-bit stream
    g2 = {}
    for name in ("SOQPSKTrellis8x1", "SOQPSKTrellis4x2", "SOQPSKTrellis4x2DiffEncoded",
                 "SimpleTrellis2", "SimpleTrellis4"):
        g2[name] = TrellisEncoder(getattr(tm, name))(bit_array)
    g2["PCMFMSymbolMapper"] = PCMFMSymbolMapper()(bit_array)
    g2["MultiHSymbolMapper"] = MultiHSymbolMapper()(bit_array)
    g2["SOQPSKPrecoder"] = SOQPSKPrecoder()(bit_array)
    # streaming: two calls on one encoder object (state persists, encoder.py:11-12)
    enc = TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)
    g2["stream_split_1001"] = np.concatenate((enc(bit_array[:1001]), enc(bit_array[1001:3000])))
    sv("g2_encode.npz", **g2)

    # G3: pulses
    g3 = {}
    for sps in (8, 10):
        g3[f"tg_{sps}"] = freq_pulse_soqpsk_tg(sps)
        g3[f"a_{sps}"] = freq_pulse_soqpsk_a(sps)
        g3[f"b_{sps}"] = freq_pulse_soqpsk_b(sps)
        g3[f"mil_{sps}"] = freq_pulse_soqpsk_mil(sps)
        g3[f"pcmfm4_{sps}"] = freq_pulse_pcmfm(sps, 4)
        g3[f"pcmfm6_{sps}"] = freq_pulse_pcmfm(sps, 6)
        g3[f"multih_{sps}"] = freq_pulse_multih_irig(sps)
        for lab, f in (("tg", freq_pulse_soqpsk_tg), ("mil", freq_pulse_soqpsk_mil)):
            r = rho_pulses(f(sps), 0.25, sps, 2)
            g3[f"rho0_{lab}_{sps}"], g3[f"rho1_{lab}_{sps}"] = r
    sv("g3_pulses.npz", **g3)

    # G4: modulated signals, 1024 bits each (irig_comparison.py:50-72 triples)
    g4 = {}
    b1k = bit_array[:1024]
    for sps in (8, 10):
        cases = {
            "pcmfm": (TrellisEncoder(tm.SimpleTrellis2)(b1k), PCMFM_NUMER / PCMFM_DENOM, freq_pulse_pcmfm(sps, 4)),
            "pcmfm6": (TrellisEncoder(tm.SimpleTrellis2)(b1k), PCMFM_NUMER / PCMFM_DENOM, freq_pulse_pcmfm(sps, 6)),
            "tg": (TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)(b1k), 0.25, freq_pulse_soqpsk_tg(sps)),
            "mil": (TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)(b1k), 0.25, freq_pulse_soqpsk_mil(sps)),
            "a": (TrellisEncoder(tm.SOQPSKTrellis4x2)(b1k), 0.25, freq_pulse_soqpsk_a(sps)),
            "b": (TrellisEncoder(tm.SOQPSKTrellis4x2)(b1k), 0.25, freq_pulse_soqpsk_b(sps)),
            "multih": (TrellisEncoder(tm.SimpleTrellis4)(b1k), MULTIH_IRIG_NUMER / MULTIH_IRIG_DENOM, freq_pulse_multih_irig(sps)),
        }
        for lab, (sy, h, g) in cases.items():
            t, s = cpm_modulate(symbols=sy, mod_index=h, pulse_filter=g, sps=sps)
            g4[f"{lab}_{sps}"] = s
            if lab == "tg" and sps == 8:
                g4["time_tg_8"] = t
    g4["bits"] = b1k
    sv("g4_modulated.npz", **g4)

    # G4b: the configs[1] / configs[2] modulators at the BASELINE frame size (N = 4096 bits, sps 8:
    # T = 32776 / 16392 samples; TG at this size is G5's tx)
    b4k = bit_array[:4096]
    g4b = {"bits": b4k}
    for lab, (sy, h, g) in {
        "pcmfm": (TrellisEncoder(tm.SimpleTrellis2)(b4k), PCMFM_NUMER / PCMFM_DENOM, freq_pulse_pcmfm(8, 4)),
        "multih": (TrellisEncoder(tm.SimpleTrellis4)(b4k), MULTIH_IRIG_NUMER / MULTIH_IRIG_DENOM, freq_pulse_multih_irig(8)),
    }.items():
        _t, s = cpm_modulate(symbols=sy, mod_index=h, pulse_filter=g, sps=8)
        g4b[f"{lab}_8"] = s
    sv("g4b_modulated_n4096.npz", **g4b)

    # G5-G8: BASELINE config 1 (first 4096 PN15 bits, sps 8, TG, 6 dB, PCG64(1))
    sps = 8
    sigma = float(np.sqrt(sps / (2 * 10 ** 0.6)))
    rng = np.random.Generator(np.random.PCG64(seed=1))
    bits = bit_array[:4096]
    r = replay_frame(bits, freq_pulse_soqpsk_tg(sps), sps, sigma, rng, {"PT": 0, "PAM": 0})
    sv("g5_config1.npz", bits=bits, symbols=r["symbols"], tx=r["tx"], noise=r["noise"], sigma=sigma,
       rho0=r["rho0"], rho1=r["rho1"],
       **{f"{k}_{f}": r[k][f] for k in ("PT", "PAM") for f in r[k]})
    print("config1", {k: (r[k]["bit_errors"], r[k]["sym_errors"], r[k]["compared"]) for k in ("PT", "PAM")})

    # non-differential + longer windows on a 512-bit frame, 8 dB, PCG64(7)
    g7 = {}
    b512 = bit_array[1000:1512]
    for diff in (True, False):
        for length in (2, 4):
            rng = np.random.Generator(np.random.PCG64(seed=7))
            rr = replay_frame(b512, freq_pulse_soqpsk_tg(8), 8, float(np.sqrt(8 / (2 * 10 ** 0.8))), rng,
                              {"PT": 0, "PAM": 0}, diff=diff, length=length)
            tag = f"d{int(diff)}_l{length}"
            g7[f"noise_{tag}"] = rr["noise"]
            for k in ("PT", "PAM"):
                g7[f"{k}_bits_{tag}"] = rr[k]["dec_bits"]
                g7[f"{k}_syms_{tag}"] = rr[k]["dec_syms"]
                g7[f"{k}_counts_{tag}"] = np.array([rr[k]["bit_errors"], rr[k]["sym_errors"], rr[k]["compared"]])
                g7[f"{k}_margin_{tag}"] = rr[k]["margin"]
                g7[f"{k}_scale_{tag}"] = rr[k]["scale"]
    g7["bits"] = b512
    sv("g7_variants.npz", **g7)

    # G11: SOQPSK-A and -B end to end (soqpsk/pulse_filters.py:53-74; 65 / 129-tap pulses, rho pulses
    # 73+65 / 137+129 taps): 1024-bit frame, sps 8, 6 dB, PCG64(11), both banks at timing offset 0
    # (probed: the noiseless replay decodes without error for offsets -2..0 on both)
    g11 = {}
    b1024 = bit_array[2000:3024]
    for lab, f in (("a", freq_pulse_soqpsk_a), ("b", freq_pulse_soqpsk_b)):
        rng = np.random.Generator(np.random.PCG64(seed=11))
        rr = replay_frame(b1024, f(8), 8, float(np.sqrt(8 / (2 * 10 ** 0.6))), rng, {"PT": 0, "PAM": 0})
        g11[f"noise_{lab}"] = rr["noise"]
        g11[f"symbols_{lab}"] = rr["symbols"]
        for k in ("PT", "PAM"):
            g11[f"{lab}_{k}_z"] = rr[k]["z"]
            g11[f"{lab}_{k}_bits"] = rr[k]["dec_bits"]
            g11[f"{lab}_{k}_syms"] = rr[k]["dec_syms"]
            g11[f"{lab}_{k}_counts"] = np.array([rr[k]["bit_errors"], rr[k]["sym_errors"], rr[k]["compared"]])
            g11[f"{lab}_{k}_margin"] = rr[k]["margin"]
            g11[f"{lab}_{k}_scale"] = rr[k]["scale"]
        print("soqpsk-" + lab, {k: tuple(g11[f"{lab}_{k}_counts"]) for k in ("PT", "PAM")})
    g11["bits"] = b1024
    g11["sigma"] = float(np.sqrt(8 / (2 * 10 ** 0.6)))
    sv("g11_soqpsk_ab.npz", **g11)

    # G12: phase trees (viz/tree.py:89-146).  tree.py imports matplotlib (absent), so its arithmetic is replayed
    # with the reference's own TrellisEncoder and cpm_modulate: sequences (:110-126), modulation + slicing
    # (:128-140), unwrap per chunk (:64-71).
    g12 = {}
    for lab, tre, h, pf in (("pcmfm", tm.SimpleTrellis2, PCMFM_NUMER / PCMFM_DENOM, freq_pulse_pcmfm(8, 4)),
                            ("multih", tm.SimpleTrellis4, MULTIH_IRIG_NUMER / MULTIH_IRIG_DENOM, freq_pulse_multih_irig(8)),
                            ("mil", tm.SOQPSKTrellis4x2DiffEncoded, 0.25, freq_pulse_soqpsk_mil(8)),
                            ("tg", tm.SOQPSKTrellis4x2DiffEncoded, 0.25, freq_pulse_soqpsk_tg(8))):
        enc = TrellisEncoder(tre)
        bps, length = enc.input_cardinality, int(pf.size // 8)
        packed_len = bps * length
        seqs = []
        for i in range(2 ** packed_len):
            enc.state = 0
            seq = [(i >> j) & 1 for j in range(packed_len)]
            seqs.append(tuple(enc.encode(np.array([0] * packed_len + seq, dtype=np.uint8))))
        seqs = sorted(set(seqs))
        plotted = np.zeros(len(seqs) * length * 8 + 1, dtype=np.complex128)
        for i, sy in enumerate(seqs):
            _t, tmp = cpm_modulate(np.array(sy, dtype=np.int8), mod_index=h, pulse_filter=pf, sps=8)
            plotted[i * length * 8:(i + 1) * length * 8] = tmp[length * 8 - 1:2 * length * 8 - 1]
        phase = np.angle(plotted)
        rows = []
        for chunk in range(phase.size // (8 * length)):
            sl = np.unwrap(phase[chunk * length * 8:(chunk + 1) * length * 8])
            rows.append(sl - sl[0])
        g12[f"{lab}_symbols"] = np.array(seqs, dtype=np.int8)
        g12[f"{lab}_phases"] = np.stack(rows)
        print("phase tree", lab, len(seqs), "sequences x", length * 8, "samples")
    sv("g12_phase_tree.npz", **g12)

    if not a.skip_example:
        # G8: the example exactly as shipped: sps 10, sigma sqrt(2)/2, MIL then TG
        # from ONE rng (soqpsk_detection.py:26,38-41,66-70,183-187)
        sps = 10
        sigma = np.sqrt(2) / 2
        rng = np.random.Generator(np.random.PCG64(seed=1))
        g8 = {}
        for lab, f, offs in (("MIL", freq_pulse_soqpsk_mil, {"PT": -1, "PAM": -3}),
                             ("TG", freq_pulse_soqpsk_tg, {"PT": -1, "PAM": 0})):
            rr = replay_frame(bit_array, f(sps), sps, sigma, rng, offs)
            g8[f"noise_crc_{lab}"] = np.array([np.sum(rr["noise"].real), np.sum(rr["noise"].imag)])
            for k in ("PT", "PAM"):
                g8[f"{lab}_{k}_counts"] = np.array([rr[k]["bit_errors"], rr[k]["sym_errors"], rr[k]["compared"]])
                g8[f"{lab}_{k}_bits"] = np.packbits(rr[k]["dec_bits"])
            print(lab, {k: (rr[k]["bit_errors"], rr[k]["sym_errors"], rr[k]["compared"]) for k in ("PT", "PAM")})
        sv("g8_example.npz", **g8)


if __name__ == "__main__":
    main()
