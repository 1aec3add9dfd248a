"""CPU: pin the oracle restatement against golden vectors produced by the
reference itself (oracle/make_golden.py).  No GPU, no /root/reference."""
import numpy as np
import pytest

from oracle import cpm_link as o
from tests.conftest import golden


# ---- G1 PRBS (reference tests/test_glfsr.py pins only periodicity; goldens pin values)
def test_masks_all_degrees():
    g = golden("g1_pn.npz")
    ours = np.array([o.pn_mask(d) for d in range(2, 65)], dtype=np.uint64)
    assert np.array_equal(ours, g["masks"])
    assert o.pn_mask(9) == 0x110 and o.pn_mask(15) == 0x6000 and o.pn_mask(23) == 0x420000


@pytest.mark.parametrize("bad", [0, 1, 65, 100, -3])
def test_mask_keyerror(bad):
    with pytest.raises(KeyError):
        o.pn_mask(bad)


@pytest.mark.parametrize("deg", [9, 15])
def test_pn_full_period(deg):
    g = golden("g1_pn.npz")
    seq = o.pn_sequence(deg)
    assert np.array_equal(np.packbits(seq), g[f"pn{deg}"])


@pytest.mark.parametrize("deg", [2, 3, 5, 7, 12, 16, 31, 32, 33, 63, 64])
def test_pn_prefix(deg):
    g = golden("g1_pn.npz")
    bits, _ = o.lfsr_bits(o.pn_mask(deg), (1 << deg) - 1, 256)
    assert np.array_equal(np.packbits(bits), g[f"pn{deg}_first256"])


def test_pn23_prefix():
    g = golden("g1_pn.npz")
    bits, _ = o.lfsr_bits(o.pn_mask(23), (1 << 23) - 1, 65536)
    assert np.array_equal(np.packbits(bits), g["pn23_first65536"])


@pytest.mark.parametrize("order", range(2, 15))
def test_pn_periodicity(order):
    """Mirror of reference tests/test_glfsr.py:6-25 (orders trimmed for CPU time)."""
    mask, st = o.pn_mask(order), (1 << order) - 1
    a, st = o.lfsr_bits(mask, st, 2**order - 1)
    b, st = o.lfsr_bits(mask, st, 2**order - 1)
    assert np.array_equal(a, b)


# ---- G2 encoders / mappers
@pytest.mark.parametrize("name", list(o.TRELLISES))
def test_encoders(name, pn15_stream):
    g = golden("g2_encode.npz")
    assert np.array_equal(o.fsm_encode(o.TRELLISES[name], pn15_stream), g[name])


def test_mappers_and_identities(pn15_stream):
    g = golden("g2_encode.npz")
    assert np.array_equal(o.map_pcmfm(pn15_stream), g["PCMFMSymbolMapper"])
    assert np.array_equal(o.map_multih(pn15_stream), g["MultiHSymbolMapper"])
    assert np.array_equal(o.soqpsk_precoder(pn15_stream), g["SOQPSKPrecoder"])
    assert np.array_equal(g["SimpleTrellis2"], g["PCMFMSymbolMapper"])
    assert np.array_equal(g["SimpleTrellis4"], g["MultiHSymbolMapper"])
    assert np.array_equal(-2 * g["SOQPSKPrecoder"], g["SOQPSKTrellis4x2"])
    assert np.array_equal(g["SOQPSKTrellis8x1"], g["SOQPSKTrellis4x2"])
    b = pn15_stream
    d = b.copy()
    for i in range(2, d.size):
        d[i] ^= d[i - 2]
    assert np.array_equal(o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2"], d), g["SOQPSKTrellis4x2DiffEncoded"])


def test_encoder_streaming(pn15_stream):
    g = golden("g2_encode.npz")
    cols = o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"]
    a, i, st = o.fsm_encode(cols, pn15_stream[:1001], return_state=True)
    b = o.fsm_encode(cols, pn15_stream[1001:3000], i0=i, state0=st)
    assert np.array_equal(np.concatenate((a, b)), g["stream_split_1001"])


def test_encoder_errors():
    with pytest.raises(ValueError):
        o.fsm_encode(o.TRELLISES["SimpleTrellis4"], np.zeros(3, dtype=np.uint8))
    with pytest.raises(ValueError):
        o.map_multih(np.zeros(5, dtype=np.uint8))


# ---- G3 pulses
@pytest.mark.parametrize("sps", [8, 10])
def test_pulses(sps):
    g = golden("g3_pulses.npz")
    cases = {
        "tg": o.pulse_soqpsk_tg(sps), "a": o.pulse_soqpsk_a(sps), "b": o.pulse_soqpsk_b(sps),
        "mil": o.pulse_soqpsk_mil(sps), "pcmfm4": o.pulse_pcmfm(sps, 4), "pcmfm6": o.pulse_pcmfm(sps, 6),
        "multih": o.pulse_multih(sps),
    }
    for k, v in cases.items():
        ref = g[f"{k}_{sps}"]
        assert v.shape == ref.shape, k
        np.testing.assert_allclose(v, ref, rtol=0, atol=1e-14, err_msg=k)
        assert abs(v.sum() - sps / 2) < 1e-12
    for lab, f in (("tg", o.pulse_soqpsk_tg), ("mil", o.pulse_soqpsk_mil)):
        r = o.rho_pulses(f(sps), 0.25, sps, 2)
        np.testing.assert_allclose(r[0], g[f"rho0_{lab}_{sps}"], rtol=0, atol=1e-14)
        np.testing.assert_allclose(r[1], g[f"rho1_{lab}_{sps}"], rtol=0, atol=1e-14)
    if sps == 8:
        assert [cases[k].size for k in ("tg", "a", "b", "mil", "pcmfm4", "pcmfm6", "multih")] == [65, 65, 129, 9, 24, 24, 25]


# ---- G4 modulators
_MOD = {
    "pcmfm": ("SimpleTrellis2", 0.7, lambda s: o.pulse_pcmfm(s, 4)),
    "pcmfm6": ("SimpleTrellis2", 0.7, lambda s: o.pulse_pcmfm(s, 6)),
    "tg": ("SOQPSKTrellis4x2DiffEncoded", 0.25, o.pulse_soqpsk_tg),
    "mil": ("SOQPSKTrellis4x2DiffEncoded", 0.25, o.pulse_soqpsk_mil),
    "a": ("SOQPSKTrellis4x2", 0.25, o.pulse_soqpsk_a),
    "b": ("SOQPSKTrellis4x2", 0.25, o.pulse_soqpsk_b),
    "multih": ("SimpleTrellis4", np.array([4, 5]) / 16, o.pulse_multih),
}


@pytest.mark.parametrize("sps", [8, 10])
@pytest.mark.parametrize("lab", list(_MOD))
def test_modulators(lab, sps):
    g = golden("g4_modulated.npz")
    tre, h, pf = _MOD[lab]
    sy = o.fsm_encode(o.TRELLISES[tre], g["bits"])
    ref = g[f"{lab}_{sps}"]
    t, s = o.cpm_modulate(sy, h, pf(sps), sps)
    assert s.shape == ref.shape
    assert np.max(np.abs(s - ref)) < 1e-12
    s2 = o.cpm_modulate_closed_form(sy, h, pf(sps), sps)
    assert np.max(np.abs(s2 - ref)) < 1e-10
    if lab == "tg" and sps == 8:
        np.testing.assert_allclose(t, g["time_tg_8"], rtol=0, atol=0)


# ---- G5-G8 config 1 chain
def test_config1_chain():
    g = golden("g5_config1.npz")
    bits = g["bits"]
    pulse = o.pulse_soqpsk_tg(8)
    for det in ("PT", "PAM"):
        r = o.soqpsk_link_frame(bits, pulse, 8, float(g["sigma"]), noise=g["noise"], detector=det,
                                fast=False, return_all=True)
        assert np.array_equal(r["symbols"], g["symbols"])
        assert np.max(np.abs(r["signal"] - g["tx"])) < 1e-12
        assert r["z"].shape == g[f"{det}_z"].shape == (4095, 3)
        assert np.max(np.abs(r["z"] - g[f"{det}_z"])) < 1e-10
        assert np.array_equal(r["dec_bits"], g[f"{det}_dec_bits"])
        assert np.array_equal(r["dec_syms"], g[f"{det}_dec_syms"])
        assert (r["bit_errors"], r["sym_errors"], r["compared"]) == (
            int(g[f"{det}_bit_errors"]), int(g[f"{det}_sym_errors"]), int(g[f"{det}_compared"]))
    assert int(g["PT_bit_errors"]) == 66 and int(g["PAM_bit_errors"]) == 32 and int(g["PAM_compared"]) == 4093


def test_fast_detector_equals_stateful():
    g = golden("g5_config1.npz")
    for det in ("PT", "PAM"):
        z = g[f"{det}_z"]
        b1, s1 = o.detect_soqpsk(z[:1500], 2, True)
        b2, s2 = o.detect_soqpsk_fast(z[:1500], True)
        assert np.array_equal(b1, b2) and np.array_equal(s1, s2)
        assert np.array_equal(b2, g[f"{det}_dec_bits"][:1500])
    zz = np.stack([g["PT_z"], g["PAM_z"]])
    bb, ss = o.detect_soqpsk_fast(zz, True)
    assert np.array_equal(bb[0], g["PT_dec_bits"]) and np.array_equal(bb[1], g["PAM_dec_bits"])
    assert np.array_equal(ss[0], g["PT_dec_syms"]) and np.array_equal(ss[1], g["PAM_dec_syms"])


@pytest.mark.parametrize("diff", [True, False])
@pytest.mark.parametrize("length", [2, 4])
def test_variants(diff, length):
    g = golden("g7_variants.npz")
    tag = f"d{int(diff)}_l{length}"
    sig = float(np.sqrt(8 / (2 * 10 ** 0.8)))
    for det in ("PT", "PAM"):
        r = o.soqpsk_link_frame(g["bits"], o.pulse_soqpsk_tg(8), 8, sig, noise=g[f"noise_{tag}"],
                                detector=det, length=length, diff=diff, fast=True, return_all=True)
        assert np.array_equal(r["dec_bits"], g[f"{det}_bits_{tag}"])
        assert np.array_equal(r["dec_syms"], g[f"{det}_syms_{tag}"])
        assert [r["bit_errors"], r["sym_errors"], r["compared"]] == list(g[f"{det}_counts_{tag}"])


def test_example_replay(pn15_stream):
    """The shipped example: sps 10, sigma sqrt(2)/2, MIL then TG from one PCG64(1)
    (soqpsk_detection.py:26,38-41,66-70,183-187).  Counts: G8."""
    g = golden("g8_example.npz")
    rng = np.random.Generator(np.random.PCG64(seed=1))
    sps, sigma = 10, np.sqrt(2) / 2
    for lab, pf, offs in (("MIL", o.pulse_soqpsk_mil, {"PT": -1, "PAM": -3}),
                          ("TG", o.pulse_soqpsk_tg, {"PT": -1, "PAM": 0})):
        T = (pn15_stream.size + 1) * sps
        noise = o.complex_awgn(sigma, T, rng)
        crc = np.array([noise.real.sum(), noise.imag.sum()])
        if not np.allclose(crc, g[f"noise_crc_{lab}"], rtol=0, atol=1e-6):
            pytest.skip("numpy PCG64 normal stream differs from the golden generator's numpy")
        for det in ("PT", "PAM"):
            r = o.soqpsk_link_frame(pn15_stream, pf(sps), sps, sigma, noise=noise, detector=det,
                                    timing_offset=offs[det], fast=True, return_all=True)
            assert [r["bit_errors"], r["sym_errors"], r["compared"]] == list(g[f"{lab}_{det}_counts"]), (lab, det)
            assert np.array_equal(np.packbits(r["dec_bits"]), g[f"{lab}_{det}_bits"])
    assert list(g["TG_PT_counts"])[::2] == [14, 32765] and list(g["TG_PAM_counts"])[::2] == [11, 32765]
    assert list(g["MIL_PT_counts"])[::2] == [61, 32765] and list(g["MIL_PAM_counts"])[::2] == [0, 32765]


# ---- G4b: configs[1] / configs[2] modulators at the BASELINE frame size (N = 4096)
@pytest.mark.parametrize("lab", ["pcmfm", "multih"])
def test_modulators_baseline_size(lab):
    g = golden("g4b_modulated_n4096.npz")
    tre, h, pf = _MOD[lab]
    sy = o.fsm_encode(o.TRELLISES[tre], g["bits"])
    ref = g[f"{lab}_8"]
    assert ref.shape == ((sy.size + 1) * 8,) and sy.size == (4096 if lab == "pcmfm" else 2048)
    assert np.max(np.abs(o.cpm_modulate(sy, h, pf(8), 8)[1] - ref)) < 1e-12
    assert np.max(np.abs(o.cpm_modulate_closed_form(sy, h, pf(8), 8) - ref)) < 1e-10


# ---- G5/G7/G11 margins: the oracle's float64 ACS margins are the ones frozen from the reference object
def test_margins_equal_the_reference_objects():
    g = golden("g5_config1.npz")
    for det in ("PT", "PAM"):
        b, s, m, sc = o.detect_soqpsk_fast(g[f"{det}_z"], True, return_margin=True)
        assert np.array_equal(b, g[f"{det}_dec_bits"])
        assert np.array_equal(m, g[f"{det}_margin"]) and np.array_equal(sc, g[f"{det}_scale"])
        assert m[0] == 0.0 and np.count_nonzero(m[1:] == 0.0) == 0  # call 0 starts from all-equal metrics
    g7 = golden("g7_variants.npz")
    for diff in (True, False):
        tag = f"d{int(diff)}_l2"
        r = o.soqpsk_link_frame(g7["bits"], o.pulse_soqpsk_tg(8), 8, 1.0, noise=g7[f"noise_{tag}"], detector="PAM",
                                diff=diff, return_all=True)
        np.testing.assert_allclose(r["margin"], g7[f"PAM_margin_{tag}"], rtol=0, atol=1e-9)


# ---- G11: SOQPSK-A / -B end to end on the reference's noise (soqpsk/pulse_filters.py:53-74)
@pytest.mark.parametrize("lab", ["a", "b"])
def test_soqpsk_a_b_chain(lab):
    g = golden("g11_soqpsk_ab.npz")
    pulse = {"a": o.pulse_soqpsk_a, "b": o.pulse_soqpsk_b}[lab](8)
    for det in ("PT", "PAM"):
        r = o.soqpsk_link_frame(g["bits"], pulse, 8, float(g["sigma"]), noise=g[f"noise_{lab}"], detector=det,
                                return_all=True)
        assert np.array_equal(r["symbols"], g[f"symbols_{lab}"])
        assert np.max(np.abs(r["z"] - g[f"{lab}_{det}_z"])) < 1e-10
        assert np.array_equal(r["dec_bits"], g[f"{lab}_{det}_bits"])
        assert np.array_equal(r["dec_syms"], g[f"{lab}_{det}_syms"])
        assert [r["bit_errors"], r["sym_errors"], r["compared"]] == list(g[f"{lab}_{det}_counts"])
        np.testing.assert_allclose(r["margin"], g[f"{lab}_{det}_margin"], rtol=0, atol=1e-9)


# ---- external anchor for the detectors the reference does not have: minimum Euclidean distance
def test_minimum_distance_reproduces_published_values():
    """oracle/cpm_dmin.py against values published for these schemes: binary CPFSK h = 0.715 -> 2.43
    (Anderson, Aulin & Sundberg 1986, the maximum for binary 1REC); ARTM CPM (M = 4, 3RC, h = 4/16, 5/16) -> 1.29
    (Perrins & Rice, IEEE Trans. AES 2007, "Reduced-complexity detectors for multi-h CPM in aeronautical
    telemetry"); IRIG PCM/FM (h = 0.7, Bessel-filtered NRZ) sits a little above CPFSK's 2.42."""
    from fractions import Fraction

    from oracle.cpm_dmin import dmin2, ebn0_db_for_ber

    rec = np.full(8, 0.5)
    assert abs(dmin2(rec, [Fraction(143, 200)], 2, 8, span=4)[0] - 2.43) < 0.01
    d_pcm = dmin2(o.pulse_pcmfm(8, 4), [Fraction(7, 10)], 2, 8, span=5)[0]
    assert 2.42 < d_pcm < 2.7
    d_artm, gam, _ = dmin2(o.pulse_multih(8), [Fraction(4, 16), Fraction(5, 16)], 4, 8, span=5)
    assert abs(d_artm - 1.29) < 0.01 and gam == (2, -4, 6, -4, 2)
    # the asymptote these distances give at BER 1e-5: 8.5 dB for PCM/FM, 11.5 dB for ARTM CPM
    assert abs(ebn0_db_for_ber(d_pcm, 1e-5) - 8.5) < 0.1 and abs(ebn0_db_for_ber(d_artm, 1e-5) - 11.5) < 0.1


# ---- diagnostics arrays (oracle/viz_ref.py) --------------------------------------------
@pytest.mark.parametrize("lab", ["pcmfm", "multih", "mil", "tg"])
def test_phase_tree_restatement_vs_reference(lab):
    """G12 = viz/tree.py:89-146 replayed with the reference's own encoder and modulator: the oracle's
    phase_tree on the oracle's modulated branches reproduces it."""
    from oracle import viz_ref as vr

    g = golden("g12_phase_tree.npz")
    hh, pf = {"pcmfm": ([0.7], o.pulse_pcmfm(8, 4)), "multih": ([4 / 16, 5 / 16], o.pulse_multih(8)),
              "mil": ([0.25], o.pulse_soqpsk_mil(8)), "tg": ([0.25], o.pulse_soqpsk_tg(8))}[lab]
    syms = g[f"{lab}_symbols"]
    length = pf.size // 8
    plotted = np.zeros(syms.shape[0] * length * 8 + 1, dtype=np.complex128)
    for i, q in enumerate(syms):
        plotted[i * length * 8:(i + 1) * length * 8] = o.cpm_modulate(q, np.array(hh), pf, 8)[1][length * 8 - 1:2 * length * 8 - 1]
    t, ph = vr.phase_tree(plotted, 8, modulo=length)
    assert ph.shape == g[f"{lab}_phases"].shape and t.size == length * 8
    assert np.max(np.abs(ph - g[f"{lab}_phases"])) < 1e-9


def test_psd_restatement_properties():
    """mlab.psd restated (parity unpinned against matplotlib, which is absent): a complex tone on a bin centre
    puts its power in that bin with the Hann leakage pattern (1, 1/4, 1/4), white noise integrates to its
    variance x sum(w^2)/sum(w)^2 x nfft, and the bps scaling multiplies power by bps."""
    from oracle import viz_ref as vr

    nfft, n = 256, 256 * 40
    k = np.arange(n)
    f, p = vr.psd(np.exp(2j * np.pi * 32 / nfft * k), 8, 1, nfft)
    pk = int(np.argmax(p))
    assert f[pk] == pytest.approx(32 / nfft * 8) and p[pk] == pytest.approx(1.0, rel=2e-2)
    assert p[pk - 1] == pytest.approx(p[pk] / 4, rel=5e-2) and p[pk + 1] == pytest.approx(p[pk] / 4, rel=5e-2)
    rng = np.random.default_rng(1)
    x = (rng.normal(size=n * 8) + 1j * rng.normal(size=n * 8)) / np.sqrt(2)
    w = np.hanning(nfft)
    _f, pn = vr.psd(x, 8, 1, nfft)
    assert np.sum(pn) == pytest.approx(nfft * np.sum(w ** 2) / np.sum(w) ** 2, rel=3e-2)
    assert np.allclose(vr.psd(x, 8, 2, nfft)[1], 2 * pn)
    lpf = vr.kaiser_fir_lpf(8, 1.0)
    assert abs(lpf.sum() - 1) < 1e-9 and np.allclose(lpf, lpf[::-1])


def test_oracle_link_reproduces_reference_ber_table_frames():
    """G10's replicates record the reference's bit-error count of EVERY frame (oracle/make_ber_table.py:
    errors_by_frame[replicate][point][frame][PT, PAM, PT@-1]).  The oracle chain fed the same PN15 frame and the same
    PCG64(1000 * EbN0 + frame + salt) noise must reproduce those counts exactly (float64 against float64): this pins
    oracle.soqpsk_link_frame -- modulator, both banks, detector, tally -- to the reference at the BER-table operating
    points, frame by frame, not only in aggregate."""
    g = golden("g10_ber_table.npz")
    if "errors_by_frame" not in g:
        pytest.skip("G10 carries no per-frame counts (single-replicate table)")
    ebf, salts = g["errors_by_frame"], [int(x) for x in g["salts"] if int(x) != 0]
    assert ebf.shape[0] == len(salts) and ebf.shape[1:] == (13, 1024, 3)
    pulse = o.pulse_soqpsk_tg(8)
    checked = 0
    for rep, salt in enumerate(salts):
        for ebn0, frame in ((0, 0), (4, 1), (4, 1023), (6, 17), (9, 500), (12, 3)):
            bits = o.pn_frame_bits(15, frame, 4096)
            sigma = float(np.sqrt(8 / (2 * 10 ** (ebn0 / 10))))
            for col, det, off in ((0, "PT", 0), (1, "PAM", 0), (2, "PT", -1)):
                rng = np.random.Generator(np.random.PCG64(seed=1000 * ebn0 + frame + salt))
                be, _se, _n = o.soqpsk_link_frame(bits, pulse, 8, sigma, rng=rng, detector=det, timing_offset=off)
                assert be == int(ebf[rep, ebn0, frame, col]), (salt, ebn0, frame, det, off, be, int(ebf[rep, ebn0, frame, col]))
                checked += 1
    assert checked == 18 * len(salts)


def test_reference_equivalent_ber_replicates_are_consistent():
    """G10b (oracle/make_ber_table_oracle.py): replicates of the BER-table experiment run through the oracle chain for
    salts the reference was not run on.  The file must say where it comes from, must have been checked against whole
    reference replicates, its salts must be disjoint from G10's, its sums must add up, and every one of its
    replicates must be statistically compatible with the reference's own replicates (same experiment, other noise):
    |replicate - pooled reference| below 5 standard errors of a replicate (design effect 2) in every cell, and the
    mean over replicates within 4 standard errors of the reference's pooled estimate."""
    import os

    from tests.conftest import GOLDEN
    if not os.path.exists(os.path.join(GOLDEN, "g10b_ber_table_oracle.npz")):
        pytest.skip("no reference-equivalent replicates in this tree")
    g, gb = golden("g10_ber_table.npz"), golden("g10b_ber_table_oracle.npz")
    assert str(gb["source"]) == "oracle" and int(gb["frames"]) == int(g["frames"]) == 1024
    assert len(gb["checked_against_g10_salts"]) >= 1 and set(map(int, gb["checked_against_g10_salts"])) <= set(map(int, g["salts"]))
    assert not set(map(int, gb["salts"])) & set(map(int, g["salts"]))
    reps = gb["errors_by_replicate"]                      # [salt][point][PT, PAM, PT@-1]
    assert reps.shape == (gb["salts"].size, 13, 3)
    for c, key in enumerate(("pt_errors", "pam_errors", "pt_m1_errors")):
        assert np.array_equal(reps[:, :, c].sum(axis=0), gb[key])
    n1 = 1024 * 4093
    assert np.all(gb["compared_by_replicate"] == n1) and np.all(gb["compared"] == reps.shape[0] * n1)
    n_ref = int(g["replicates"])
    for c, key in enumerate(("pt_errors", "pam_errors", "pt_m1_errors")):
        p_ref = g[key] / (n_ref * n1)
        se1 = np.sqrt(2.0 * p_ref * (1 - p_ref) / n1)     # one replicate, design effect 2
        dev = (reps[:, :, c] / n1 - p_ref) / np.sqrt(se1 ** 2 * (1 + 1 / n_ref))
        assert np.max(np.abs(dev)) < 5.0, (key, float(np.max(np.abs(dev))))
        zmean = (reps[:, :, c].mean(axis=0) / n1 - p_ref) / np.sqrt(se1 ** 2 * (1 / reps.shape[0] + 1 / n_ref))
        assert np.max(np.abs(zmean)) < 4.0, (key, np.round(zmean, 2).tolist())
