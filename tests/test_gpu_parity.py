"""GPU parity tests: the CUDA path (through the Python mirror of the reference API,
i.e. through the C ABI of libwfb200.so) against the oracle and the golden vectors the
reference produced.  Integer work is bit-exact; float tolerances are written at each
assert.  Run on the B200 box: python -m pytest tests -m gpu."""
import numpy as np
import pytest
import torch

from oracle import cpm_link as o
from tests.conftest import golden

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import waveforms_b200 as wf  # noqa: F401
    from waveforms_b200 import _ffi, bert, mfbank
    from waveforms_b200.cpm import modulate as wmod
    from waveforms_b200.cpm.multih import MultiHSymbolMapper, freq_pulse_multih_irig
    from waveforms_b200.cpm.pcmfm import PCMFMSymbolMapper, freq_pulse_pcmfm
    from waveforms_b200.cpm.soqpsk import (SOQPSKPrecoder, freq_pulse_soqpsk_a, freq_pulse_soqpsk_b,
                                          freq_pulse_soqpsk_mil, freq_pulse_soqpsk_tg)
    from waveforms_b200.cpm.trellis import model as tm
    from waveforms_b200.cpm.trellis.encoder import TrellisEncoder
    from waveforms_b200.glfsr import GLFSR, PNSequence
    from waveforms_b200.linksim import LinkConfig, LinkSimulator
    from waveforms_b200.noise import PhiloxRNG, add_awgn, generate_complex_awgn
    from waveforms_b200.viterbi import SOQPSKTrellisDetector

DEV = "cuda:0"


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(DEV)


REL_TIE = 1e-5  # a float32 decision may differ from the float64 one only where margin < 1e-5


def assert_fp32_flips_are_near_ties(dec32, dec64, margin, scale, window=1, what=""):
    """SURVEY 8(c) acceptance for the float32 path: every decision that differs from the float64
    detector's must sit on a step whose float64 decision margin (frozen from the reference object by
    oracle/make_golden.py:acs_margin, or computed by the oracle the same way) is below REL_TIE x the
    largest candidate metric of that step -- i.e. a genuine near-tie, not an ordering bug.  ``window``:
    steps [k, k + window) can influence decision k (1 for the SOQPSK detector, the survivor memory for the
    block-traceback CPM detectors).  Prints the relative-margin histogram."""
    dec32, dec64 = np.asarray(dec32), np.asarray(dec64)
    rel = np.asarray(margin, dtype=np.float64) / np.maximum(np.asarray(scale, dtype=np.float64), 1.0)
    hist = np.histogram(np.log10(np.maximum(rel, 1e-12)), bins=[-12, -7, -6, -5, -4, -3, -2, -1, 1])[0]
    flips = np.flatnonzero(dec32 != dec64)
    print(f"{what}: {flips.size} float32 flips of {dec64.size}; relative margin histogram "
          f"(<1e-7, 1e-7..1e-6, ..1e-5, ..1e-4, ..1e-3, ..1e-2, ..1e-1, >=1e-1) = {hist.tolist()}")
    for k in flips:
        m = float(rel[k:k + window].min())
        assert m < REL_TIE, f"{what}: decision {k} differs but the smallest float64 margin in reach is {m:.3e}"
    # and near-ties are rare enough that the check has teeth (a wide trellis makes hundreds of comparisons
    # per step, so a few per cent of the steps see one near-tie among them)
    assert np.count_nonzero(rel < REL_TIE) <= max(4, dec64.size // 20), (what, hist.tolist())
    return flips.size


def test_library_is_native_and_loaded():
    lib = _ffi.lib()
    assert lib.wf_abi_version() == _ffi.ABI_VERSION
    assert lib.wf_device_check() == 0, lib.wf_last_error_string()
    with open("/proc/self/maps") as f:
        assert "libwfb200.so" in f.read()


# ------------------------------------------------------------------ A1 PRBS
@pytest.mark.parametrize("deg", [9, 15])
def test_pn_full_period(deg):
    g = golden("g1_pn.npz")
    seq = PNSequence(deg).generate_sequence()
    assert isinstance(seq, list) and len(seq) == 2 ** deg - 1
    assert np.array_equal(np.packbits(np.array(seq, dtype=np.uint8)), g[f"pn{deg}"])


@pytest.mark.parametrize("deg", [2, 3, 5, 7, 12, 16, 31, 32, 33, 63, 64])
def test_pn_prefix_all_widths(deg):
    g = golden("g1_pn.npz")
    bits = PNSequence(deg).next_bits(256).cpu().numpy()
    assert np.array_equal(np.packbits(bits), g[f"pn{deg}_first256"])


def test_pn23_prefix_and_next_bit():
    g = golden("g1_pn.npz")
    p = PNSequence(23)
    bits = p.next_bits(65536).cpu().numpy()
    assert np.array_equal(np.packbits(bits), g["pn23_first65536"])
    # next_bit() continues the same stream and keeps .state exact
    q = PNSequence(23)
    first = [q.next_bit() for _ in range(100)]
    assert first == list(np.unpackbits(g["pn23_first65536"])[:100])
    _, st = o.lfsr_bits(o.pn_mask(23), (1 << 23) - 1, 100)
    assert q.state == st
    r = GLFSR(o.pn_mask(9), 0x1FF)
    assert [r.next_bit() for _ in range(40)] == list(o.pn_sequence(9)[:40])


@pytest.mark.parametrize("order", range(2, 23))
def test_pn_periodicity(order):
    """Mirror of the reference's only test, tests/test_glfsr.py:6-25."""
    pn = PNSequence(order)
    a = pn.generate_sequence()
    b = pn.generate_sequence()
    assert a == b


def test_pn_frames_jump_ahead():
    """Frame f = stream bits [f*N, (f+1)*N) mod period, for far-away frames too."""
    pn = PNSequence(15)
    for first, B, N in ((0, 7, 4096), (123456789, 5, 1000), (2 ** 40 + 3, 3, 333)):
        fr = pn.frames(first, B, N).cpu().numpy()
        for b in range(B):
            assert np.array_equal(fr[b], o.pn_frame_bits(15, first + b, N)), (first, b)
        pk = pn.frames(first, B, N, packed=True).cpu().numpy().view(np.uint32)
        ref = np.packbits(fr, axis=1, bitorder="little")
        ref = np.concatenate((ref, np.zeros((B, (-ref.shape[1]) % 4), np.uint8)), axis=1).view(np.uint32)
        assert np.array_equal(pk, ref)


# ------------------------------------------------------------------ A2-A4 encoders
@pytest.mark.parametrize("name", ["SOQPSKTrellis8x1", "SOQPSKTrellis4x2", "SOQPSKTrellis4x2DiffEncoded",
                                  "SimpleTrellis2", "SimpleTrellis4"])
def test_trellis_encoders(name, pn15_stream):
    g = golden("g2_encode.npz")
    enc = TrellisEncoder(getattr(tm, name))
    out = enc(cu(pn15_stream)).cpu().numpy()
    assert out.dtype == np.int8 and np.array_equal(out, g[name])


def test_mappers_precoder_and_streaming(pn15_stream):
    g = golden("g2_encode.npz")
    b = cu(pn15_stream)
    assert np.array_equal(PCMFMSymbolMapper()(b).cpu().numpy(), g["PCMFMSymbolMapper"])
    assert np.array_equal(MultiHSymbolMapper()(b).cpu().numpy(), g["MultiHSymbolMapper"])
    assert np.array_equal(SOQPSKPrecoder()(b).cpu().numpy(), g["SOQPSKPrecoder"])
    enc = TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)
    two = torch.cat((enc(b[:1001]), enc(b[1001:3000]))).cpu().numpy()
    assert np.array_equal(two, g["stream_split_1001"])
    pre = SOQPSKPrecoder()
    parts = torch.cat((pre(b[:501]), pre(b[501:1500]), pre(b[1500:1503]))).cpu().numpy()
    assert np.array_equal(parts, g["SOQPSKPrecoder"][:1503])


def test_encoder_batch_ragged_and_errors(pn15_stream):
    cols = o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"]
    for n in (1, 2, 31, 33, 1000):
        frames = np.stack([pn15_stream[k * 1000:k * 1000 + n] for k in range(5)])
        out = TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)(cu(frames)).cpu().numpy()
        for k in range(5):
            assert np.array_equal(out[k], o.fsm_encode(cols, frames[k])), (n, k)
    with pytest.raises(ValueError, match="multiple of FSM cardinality"):
        TrellisEncoder(tm.SimpleTrellis4)(cu(np.zeros(3, np.uint8)))
    with pytest.raises(ValueError, match="Odd length"):
        MultiHSymbolMapper()(cu(np.zeros(5, np.uint8)))
    with pytest.raises(TypeError):
        TrellisEncoder(tm.SimpleTrellis2)(torch.zeros(8, dtype=torch.uint8))  # CPU tensor: no fallback
    with pytest.raises(KeyError):
        PNSequence(65)


# ------------------------------------------------------------------ A6 modulators
_MOD = {
    "pcmfm": ("SimpleTrellis2", 0.7, lambda s: freq_pulse_pcmfm(s, 4)),
    "pcmfm6": ("SimpleTrellis2", 0.7, lambda s: freq_pulse_pcmfm(s, 6)),
    "tg": ("SOQPSKTrellis4x2DiffEncoded", 0.25, lambda s: freq_pulse_soqpsk_tg(s)),
    "mil": ("SOQPSKTrellis4x2DiffEncoded", 0.25, lambda s: freq_pulse_soqpsk_mil(s)),
    "a": ("SOQPSKTrellis4x2", 0.25, lambda s: freq_pulse_soqpsk_a(s)),
    "b": ("SOQPSKTrellis4x2", 0.25, lambda s: freq_pulse_soqpsk_b(s)),
    "multih": ("SimpleTrellis4", np.array([4, 5]) / 16, lambda s: freq_pulse_multih_irig(s)),
}


@pytest.mark.parametrize("sps", [8, 10])
@pytest.mark.parametrize("lab", list(_MOD))
def test_modulators_vs_reference(lab, sps):
    """north_star: modulated baseband within 1e-5 of the reference's float64 (|s| = 1, so
    absolute = relative); the float64 instantiation agrees to 1e-10."""
    g = golden("g4_modulated.npz")
    tre, h, pf = _MOD[lab]
    sy = cu(o.fsm_encode(o.TRELLISES[tre], g["bits"]))
    ref = g[f"{lab}_{sps}"]
    t, s = wmod.cpm_modulate(sy, h, pf(sps), sps)
    assert s.dtype == torch.complex64 and s.shape == ref.shape
    assert np.max(np.abs(s.cpu().numpy() - ref)) < 1e-5
    _, s64 = wmod.cpm_modulate(sy, h, pf(sps), sps, dtype=torch.complex128)
    assert np.max(np.abs(s64.cpu().numpy() - ref)) < 1e-10
    if lab == "tg" and sps == 8:
        np.testing.assert_allclose(t.cpu().numpy(), g["time_tg_8"], rtol=0, atol=0)


@pytest.mark.parametrize("lab", ["pcmfm", "multih"])
def test_modulators_vs_reference_at_baseline_size(lab):
    """BASELINE configs[1] / configs[2] modulators at N = 4096 bits (G4b, produced by the reference):
    1e-5 for the float32 throughput kernels (PCM/FM takes the ODD fast path), 1e-10 for float64."""
    g = golden("g4b_modulated_n4096.npz")
    tre, h, pf = _MOD[lab]
    sy = cu(o.fsm_encode(o.TRELLISES[tre], g["bits"]))
    ref = g[f"{lab}_8"]
    _, s = wmod.cpm_modulate(sy, h, pf(8), 8)
    assert s.shape == ref.shape and np.max(np.abs(s.cpu().numpy() - ref)) < 1e-5
    _, s64 = wmod.cpm_modulate(sy, h, pf(8), 8, dtype=torch.complex128)
    assert np.max(np.abs(s64.cpu().numpy() - ref)) < 1e-10
    # the same frame inside a batch (other rows = shifted copies) is unchanged
    syb = torch.stack([sy, torch.roll(sy, 5), torch.roll(sy, 11)])
    assert torch.equal(wmod.modulate_frames(syb, h, pf(8), 8)[0], s)


def test_modulator_long_frames_batch():
    """Several 1024-symbol chunks per frame, batch axis, no drift of the fixed-point phase."""
    bits = np.unpackbits(golden("g1_pn.npz")["pn15"])
    frames = np.stack([np.roll(bits, -k * 4099)[:20000] for k in range(3)])
    sy = np.stack([o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], f) for f in frames])
    pulse = freq_pulse_soqpsk_tg(8)
    _, s = wmod.cpm_modulate(cu(sy), 0.25, pulse, 8)
    for k in range(3):
        ref = o.cpm_modulate_closed_form(sy[k], 0.25, pulse, 8)
        assert np.max(np.abs(s[k].cpu().numpy() - ref)) < 1e-5


@pytest.mark.parametrize("lab", ["tg", "multih", "mil", "pcmfm"])
def test_modulator_degenerate_and_segment_boundary_lengths(lab):
    """The sps-8 fast path scans a frame in segments of 5120 symbol intervals, 20 symbols per thread:
    one-symbol frames, lengths around the per-thread, per-step and per-segment boundaries, with the
    fused noise equal to the stand-alone AWGN kernel's on every length."""
    rng = np.random.default_rng(21)
    tre, h, pf = _MOD[lab]
    alpha = {"multih": (-3, -1, 1, 3), "pcmfm": (-1, 1)}.get(lab, (-2, 0, 2))
    pulse = pf(8)
    for S in (1, 2, 7, 8, 9, 20, 21, 1023, 1025, 5099, 5116, 5119, 5120, 5121, 10241):
        B = 3 if S < 2000 else 1
        sy = rng.choice(np.array(alpha, np.int8), size=(B, S)).astype(np.int8)
        clean = wmod.modulate_frames(cu(sy), h, pulse, 8).contiguous()
        for b in range(B):
            ref = o.cpm_modulate_closed_form(sy[b], h, pulse, 8)
            assert np.max(np.abs(clean[b].cpu().numpy() - ref)) < 1e-5, (lab, S, b)
        sig = torch.full((B,), 0.8, dtype=torch.float32, device=DEV)
        sep = add_awgn(clean.clone(), sig, PhiloxRNG(5), first_frame=9)
        fused = wmod.modulate_frames(cu(sy), h, pulse, 8, sigma=sig, seed=5, first_frame=9)
        assert torch.max(torch.abs(sep - fused)).item() < 2e-6, (lab, S)


def test_frequency_and_phase_modulate():
    rng = np.random.default_rng(3)
    f = rng.normal(size=3000) * 0.3
    out = wmod.frequency_modulate(cu(f), 8, 0.4).cpu().numpy()
    revs, ph = 0.0, np.zeros(3000)
    for i in range(3000):
        revs = (revs + f[i]) % 8
        ph[i] = revs * 2 * np.pi / 8 + 0.4
    assert np.max(np.abs(out - np.exp(1j * ph))) < 1e-11
    pm = wmod.phase_modulate(cu(f), 1.7).cpu().numpy()
    assert np.max(np.abs(pm - np.exp(1j * 1.7 * f))) < 1e-12


# ------------------------------------------------------------------ A7 noise
def test_awgn_statistics_and_sharding_independence():
    rng = PhiloxRNG(seed=1234)
    x = generate_complex_awgn(0.5, 1 << 20, rng)
    v = torch.view_as_real(x).cpu().numpy().astype(np.float64) / 0.5
    assert x.dtype == torch.complex64 and x.shape == (1 << 20,)
    n = v.size
    assert abs(v.mean()) < 5 / np.sqrt(n)
    assert abs(v.var() - 1) < 5 * np.sqrt(2 / n)
    assert abs(np.mean(v ** 4) - 3) < 5 * np.sqrt(96 / n)
    assert abs(np.mean(v[:, 0] * v[:, 1])) < 5 / np.sqrt(n / 2)       # re/im uncorrelated
    assert abs(np.mean(v[:-1, 0] * v[1:, 0])) < 5 / np.sqrt(n / 2)    # white
    # tail: P(|x| > 4) = 6.33e-5 per real component
    tail = np.mean(np.abs(v) > 4.0)
    assert 0.7 * 6.33e-5 < tail < 1.3 * 6.33e-5
    # frames (seed, frame id) reproduce regardless of batching
    a = generate_complex_awgn(1.0, 4096, PhiloxRNG(7, first_frame=0), batch=8)
    b = generate_complex_awgn(1.0, 4096, PhiloxRNG(7, first_frame=5), batch=3)
    assert torch.equal(a[5:], b)
    assert not torch.equal(a[0], a[1])


def test_awgn_distribution_large_sample():
    """2^27 real samples: moments, tail probabilities on both sides of the 16-bit radius grid edge
    (4.71 sigma, where the tail refinement takes over), radius and angle laws."""
    from math import erfc, exp, sqrt
    B, T = 2048, 32768
    x = generate_complex_awgn(1.0, T, PhiloxRNG(seed=2024, first_frame=123), batch=B)
    v = torch.view_as_real(x).double()
    n = v.numel()
    assert abs(v.mean().item()) < 5 / sqrt(n)
    assert abs((v * v).mean().item() - 1) < 5 * sqrt(2 / n)
    assert abs((v ** 4).mean().item() - 3) < 5 * sqrt(96 / n)
    assert abs((v ** 6).mean().item() - 15) < 5 * sqrt(10170 / n)
    assert abs((v[..., 0] * v[..., 1]).mean().item()) < 5 / sqrt(n / 2)
    a = v.abs()
    for t in (1.0, 2.0, 3.0, 4.0, 4.5, 5.0):
        p = erfc(t / sqrt(2))
        got = (a > t).double().mean().item()
        assert abs(got - p) < 5 * sqrt(p / n) + 1e-12, (t, got, p)
    r2 = (v * v).sum(-1)  # chi^2 with 2 dof: P(r^2 > s) = exp(-s/2)
    m = r2.numel()
    for s_ in (0.01, 0.5, 2.0, 8.0, 16.0, 22.0, 26.0):
        p = exp(-s_ / 2)
        got = (r2 > s_).double().mean().item()
        assert abs(got - p) < 5 * sqrt(p * (1 - p) / m) + 1e-12, (s_, got, p)
    ang = torch.atan2(v[..., 1], v[..., 0])
    hist = torch.histc(ang.float(), bins=64, min=-np.pi, max=np.pi).double()
    chi2 = (((hist - m / 64) ** 2) / (m / 64)).sum().item()
    assert chi2 < 63 + 5 * sqrt(2 * 63), chi2
    # neighbouring samples (which share a Philox call with the pair 4 pairs on) are uncorrelated
    f = v[..., 0]
    for lag in (1, 2, 8, 9):
        c = (f[:, :-lag] * f[:, lag:]).mean().item()
        assert abs(c) < 5 / sqrt(f[:, lag:].numel()), (lag, c)


def test_fused_noise_equals_separate_add():
    g = golden("g4_modulated.npz")
    sy = cu(o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], g["bits"]))[None].repeat(4, 1)
    pulse = freq_pulse_soqpsk_tg(8)
    sig = torch.tensor([0.3, 0.7, 1.0, 2.0], dtype=torch.float32, device=DEV)
    pt = torch.tensor([0, 1, 2, 3], dtype=torch.int32, device=DEV)
    clean = wmod.modulate_frames(sy, [0.25], pulse, 8, derotate=-np.pi / 4).contiguous()
    sep = add_awgn(clean.clone(), sig, PhiloxRNG(99), point=pt, first_frame=5)
    fused = wmod.modulate_frames(sy, [0.25], pulse, 8, derotate=-np.pi / 4, sigma=sig, point=pt, seed=99, first_frame=5)
    assert torch.max(torch.abs(sep - fused)).item() < 1e-6
    noise = (sep - clean)
    assert abs(torch.view_as_real(noise[3]).std().item() / 2.0 - 1) < 0.05


# ------------------------------------------------------------------ A8/A9 banks, A10 detector, A13 count
def _rx_config1():
    g = golden("g5_config1.npz")
    return g, g["tx"] * np.exp(-1j * np.pi / 4) + g["noise"]


@pytest.mark.parametrize("det", ["PT", "PAM"])
def test_banks_vs_reference(det):
    g, rx = _rx_config1()
    pulse = freq_pulse_soqpsk_tg(8)
    bank = mfbank.pt_bank if det == "PT" else mfbank.pam_bank
    z64 = bank(cu(rx), pulse, 0.25, 8).cpu().numpy()
    assert z64.shape == (4095, 3)
    assert np.max(np.abs(z64 - g[f"{det}_z"])) < 1e-9
    z32 = bank(cu(rx, torch.complex64), pulse, 0.25, 8).cpu().numpy()
    # float32 accumulation of <= 73 taps on |rx| ~ 1.7: 1e-4 absolute on outputs of magnitude ~ 8
    assert np.max(np.abs(z32 - g[f"{det}_z"])) < 1e-4


def test_generic_bank_shape():
    rng = np.random.default_rng(0)
    taps = rng.normal(size=(5, 13)) + 1j * rng.normal(size=(5, 13))
    rr = rng.normal(size=(2, 1000)) + 1j * rng.normal(size=(2, 1000))
    z = mfbank.filter_bank(cu(rr), taps, sps=8, first_instant=3, n_calls=100).cpu().numpy()
    for b in range(2):
        full = np.stack([np.convolve(rr[b], t, "same") for t in taps])
        assert np.max(np.abs(z[b] - full[:, 3 + 8 * np.arange(100)].T)) < 1e-9


@pytest.mark.parametrize("det", ["PT", "PAM"])
def test_detector_bit_exact_on_reference_noise(det):
    """north_star: bit-exact Viterbi decisions when fed the reference's own noise samples.
    float64 instantiation: every one of the 4095 decisions and the counts of G8."""
    g, rx = _rx_config1()
    pulse = freq_pulse_soqpsk_tg(8)
    bank = mfbank.pt_bank if det == "PT" else mfbank.pam_bank
    z = bank(cu(rx)[None], pulse, 0.25, 8)
    tx_bits = cu(g["bits"])[None]
    tx_syms = cu(g["symbols"])[None]
    pk = np.packbits(g["bits"], bitorder="little").view(np.uint32).view(np.int32)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    db, ds = SOQPSKTrellisDetector().detect(z, tx_packed=cu(pk)[None], tx_symbols=tx_syms, counters=counters)
    assert np.array_equal(db[0].cpu().numpy(), g[f"{det}_dec_bits"])
    assert np.array_equal(ds[0].cpu().numpy(), g[f"{det}_dec_syms"])
    want = [int(g[f"{det}_bit_errors"]), int(g[f"{det}_sym_errors"]), int(g[f"{det}_compared"])]
    assert counters[0].tolist() == want
    assert bert.count_errors(db, tx_bits, 2, dec_symbols=ds, ref_symbols=tx_syms) == tuple(want)
    # float32 throughput path, same input: decisions may differ only where the reference's own float64
    # ACS margin (G5 *_margin / *_scale) is at float32 resolution
    z32 = bank(cu(rx, torch.complex64)[None], pulse, 0.25, 8)
    db32, ds32 = SOQPSKTrellisDetector().detect(z32)
    assert_fp32_flips_are_near_ties(db32[0].cpu().numpy(), g[f"{det}_dec_bits"], g[f"{det}_margin"], g[f"{det}_scale"],
                                    what=f"config-1 {det} bits")
    assert_fp32_flips_are_near_ties(ds32[0].cpu().numpy(), g[f"{det}_dec_syms"], g[f"{det}_margin"], g[f"{det}_scale"],
                                    what=f"config-1 {det} symbols")


@pytest.mark.parametrize("diff", [True, False])
@pytest.mark.parametrize("length", [2, 4])
def test_detector_variants(diff, length):
    g = golden("g7_variants.npz")
    tag = f"d{int(diff)}_l{length}"
    pulse = freq_pulse_soqpsk_tg(8)
    sig = float(np.sqrt(8 / (2 * 10 ** 0.8)))
    for det, bank in (("PT", mfbank.pt_bank), ("PAM", mfbank.pam_bank)):
        r = o.soqpsk_link_frame(g["bits"], pulse, 8, sig, noise=g[f"noise_{tag}"], detector=det, length=length,
                                diff=diff, return_all=True)
        z = bank(cu(r["rx"])[None], pulse, 0.25, 8, length=length)
        db, ds = SOQPSKTrellisDetector(length, differantial_encoding=diff).detect(z)
        assert np.array_equal(db[0].cpu().numpy(), g[f"{det}_bits_{tag}"])
        assert np.array_equal(ds[0].cpu().numpy(), g[f"{det}_syms_{tag}"])
        z32 = bank(cu(r["rx"], torch.complex64)[None], pulse, 0.25, 8, length=length)
        db32, _ = SOQPSKTrellisDetector(length, differantial_encoding=diff).detect(z32)
        assert_fp32_flips_are_near_ties(db32[0].cpu().numpy(), g[f"{det}_bits_{tag}"], g[f"{det}_margin_{tag}"],
                                        g[f"{det}_scale_{tag}"], what=f"G7 {det} {tag}")


@pytest.mark.parametrize("lab", ["a", "b"])
def test_soqpsk_a_b_end_to_end_on_reference_noise(lab):
    """SOQPSK-A / -B (soqpsk/pulse_filters.py:53-74; B has a 129-tap pulse and 137/129-tap rho
    pulses) through modulator -> bank -> detector -> count, fed the REFERENCE's own noise (G11): float64
    instantiation identical on every decision and count, float32 flips only at near-ties."""
    g = golden("g11_soqpsk_ab.npz")
    pf = {"a": freq_pulse_soqpsk_a, "b": freq_pulse_soqpsk_b}[lab]
    pulse = pf(8)
    syms = TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)(cu(g["bits"]))
    assert np.array_equal(syms.cpu().numpy(), g[f"symbols_{lab}"])
    _, s = wmod.cpm_modulate(syms, 0.25, pulse, 8, dtype=torch.complex128)
    rx = (s * np.exp(-1j * np.pi / 4) + cu(g[f"noise_{lab}"]))[None]
    pk = cu(np.packbits(g["bits"], bitorder="little").view(np.uint32).view(np.int32))[None]
    for det, bank in (("PT", mfbank.pt_bank), ("PAM", mfbank.pam_bank)):
        z = bank(rx, pulse, 0.25, 8)
        assert np.max(np.abs(z[0].cpu().numpy() - g[f"{lab}_{det}_z"])) < 1e-9
        counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
        db, ds = SOQPSKTrellisDetector().detect(z, tx_packed=pk, tx_symbols=syms[None], counters=counters)
        assert np.array_equal(db[0].cpu().numpy(), g[f"{lab}_{det}_bits"])
        assert np.array_equal(ds[0].cpu().numpy(), g[f"{lab}_{det}_syms"])
        assert counters[0].tolist() == list(g[f"{lab}_{det}_counts"]), (lab, det)
        # float32 chain (float32 modulator included: the throughput path end to end)
        _, s32 = wmod.cpm_modulate(syms, 0.25, pulse, 8)
        rx32 = (s32 * complex(np.exp(-1j * np.pi / 4)) + cu(g[f"noise_{lab}"], torch.complex64))[None].contiguous()
        db32, _ = SOQPSKTrellisDetector().detect(bank(rx32, pulse, 0.25, 8))
        assert_fp32_flips_are_near_ties(db32[0].cpu().numpy(), g[f"{lab}_{det}_bits"], g[f"{lab}_{det}_margin"],
                                        g[f"{lab}_{det}_scale"], what=f"SOQPSK-{lab.upper()} {det}")


@pytest.mark.parametrize("sps,off", [(8, 0), (10, 0), (10, -3), (8, -2)])
def test_mil_pam_bank_compiled_instances_vs_float64(sps, off):
    """SOQPSK-MIL's PAM bank (17 taps at sps 8, 21 at sps 10; SURVEY 8(f)1: MIL needs timing offset -3 at sps 10)
    has compiled register-tiled float32 instances for an even tile origin; an odd origin (offset -3) takes the
    run-time-shape kernel.  Both against the float64 bank on the same noisy frames: float32 accumulation of <= 21
    taps on |rx| ~ 1.7 stays below 1e-4 absolute, and the float64 bank is the reference's np.convolve to 1e-9."""
    rng = np.random.default_rng(sps + off)
    pulse = freq_pulse_soqpsk_mil(sps)
    bits = rng.integers(0, 2, size=(3, 700)).astype(np.uint8)
    syms = np.stack([o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], b) for b in bits])
    sig = np.stack([o.cpm_modulate_closed_form(q, 0.25, pulse, sps) for q in syms]) * np.exp(-1j * np.pi / 4)
    rx = sig + 0.8 * (rng.normal(size=sig.shape) + 1j * rng.normal(size=sig.shape))
    z64 = mfbank.pam_bank(cu(rx), pulse, 0.25, sps, timing_offset=off).cpu().numpy()
    z32 = mfbank.pam_bank(cu(rx, torch.complex64), pulse, 0.25, sps, timing_offset=off).cpu().numpy()
    assert z64.shape == z32.shape and z64.shape[0] == 3 and z64.shape[2] == 3
    assert np.max(np.abs(z32 - z64)) < 1e-4
    first, calls = mfbank.symbol_instants(rx.shape[1], sps, off, 2)
    pam = o.pam_bank(rx[1], pulse, 0.25, sps)  # soqpsk_detection.py:159-173 at full rate (np.convolve "same")
    assert np.max(np.abs(z64[1] - pam[:, first + sps * np.arange(calls)].T)) < 1e-9


@pytest.mark.parametrize("lab", ["a", "b", "mil"])
def test_soqpsk_a_b_link_simulator_noiseless_and_monotone(lab):
    """LinkSimulator on SOQPSK-A / -B / -MIL at the BASELINE frame size: noiseless frames decode without error
    (the reference replay does, offsets -2..0), BER falls with Eb/N0."""
    for det in ("PAM", "PT"):
        sim = LinkSimulator(LinkConfig(waveform=f"soqpsk-{lab}", detector=det), device=DEV, max_batch=256)
        counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
        sim.run_batch(3, torch.zeros(256, dtype=torch.float32, device=DEV), None, counters)
        assert counters[0].tolist()[::2] == [0, 256 * 4093], (lab, det)
        got = sim.run([2.0, 6.0, 10.0], 128)
        ber = got[:, 0] / got[:, 2]
        # 10 dB: A / B below 3e-3; MIL's PT detector reads 3.0e-3 at sps 8 / offset 0 (the shipped example's MIL replay, G8, reads 1.9e-3 at sps 10 with its own offset)
        assert ber[0] > ber[1] > ber[2] > 0 and ber[2] < (6e-3 if lab == "mil" else 3e-3), (lab, det, ber)


def test_detector_streaming_and_odd_windows():
    g = golden("g5_config1.npz")
    z = g["PAM_z"][:301]
    for length in (1, 2, 3, 6, 16):
        rb, rs = o.detect_soqpsk(z, length, True)
        det = SOQPSKTrellisDetector(length)
        parts = [det.iterate(cu(z[a:b])) for a, b in ((0, 101), (101, 102), (102, 301))]
        assert np.array_equal(torch.cat([p[0] for p in parts]).cpu().numpy(), rb), length
        assert np.array_equal(torch.cat([p[1] for p in parts]).cpu().numpy(), rs), length
    det = SOQPSKTrellisDetector()
    rb, rs = o.detect_soqpsk(z[:20], 2, True)
    for c in range(20):
        b, s = det.iteration(cu(z[c]))
        assert b[0].item() == rb[c] and s[0].item() == rs[c]


def test_example_replay_counts(pn15_stream):
    """The shipped example (sps 10, sigma sqrt(2)/2, MIL then TG from one PCG64(1)): the CUDA
    chain fed the same noise reproduces G8's error counts (TG PT 14 / PAM 11 of 32765, ...)."""
    g = golden("g8_example.npz")
    rng = np.random.Generator(np.random.PCG64(seed=1))
    sps, sigma = 10, np.sqrt(2) / 2
    bits = cu(pn15_stream)
    syms = TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)(bits)
    pk = cu(np.packbits(pn15_stream, bitorder="little").view(np.uint32).view(np.int32))[None]
    for lab, pf, offs in (("MIL", freq_pulse_soqpsk_mil, {"PT": -1, "PAM": -3}),
                          ("TG", freq_pulse_soqpsk_tg, {"PT": -1, "PAM": 0})):
        T = (pn15_stream.size + 1) * sps
        noise = o.complex_awgn(sigma, T, rng)
        crc = np.array([noise.real.sum(), noise.imag.sum()])
        if not np.allclose(crc, g[f"noise_crc_{lab}"], rtol=0, atol=1e-6):
            pytest.skip("numpy PCG64 normal stream differs from the golden generator's numpy")
        _, s = wmod.cpm_modulate(syms, 0.25, pf(sps), sps, dtype=torch.complex128)
        rx = (s * np.exp(-1j * np.pi / 4) + cu(noise))[None]
        for det, bank in (("PT", mfbank.pt_bank), ("PAM", mfbank.pam_bank)):
            z = bank(rx, pf(sps), 0.25, sps, timing_offset=offs[det])
            counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
            db, _ = SOQPSKTrellisDetector().detect(z, tx_packed=pk, tx_symbols=syms[None], counters=counters)
            assert counters[0].tolist() == list(g[f"{lab}_{det}_counts"]), (lab, det)
            assert np.array_equal(np.packbits(db[0].cpu().numpy()), g[f"{lab}_{det}_bits"])


# ------------------------------------------------------------------ whole chain, full size
def test_link_noiseless_round_trip_full_batch():
    """Size-independent property at BASELINE size (N=4096, sps=8): with sigma = 0 every frame
    decodes without error, for both banks."""
    for det in ("PAM", "PT"):
        sim = LinkSimulator(LinkConfig(detector=det), device=DEV, max_batch=2048)
        counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
        sim.run_batch(1000, torch.zeros(2048, dtype=torch.float32, device=DEV), None, counters)
        be, se, n = counters[0].tolist()
        assert (be, n) == (0, 2048 * 4093), det


def test_raw_bank_detection_equals_combined_bank_detection():
    """PAM link, two data paths: the bank stores z[.., 3] and wf_viterbi_soqpsk reads it, or the bank
    stores its two raw outputs and wf_viterbi_soqpsk_combined applies the pseudo-symbol combine.
    Counters must be identical at every Eb/N0 point (same fused operations, same order)."""
    eb = [0.0, 4.0, 8.0]
    a = LinkSimulator(LinkConfig(detector="PAM", n_bits=2048), device=DEV, max_batch=96)
    b = LinkSimulator(LinkConfig(detector="PAM", n_bits=2048, extra={"combine_in_bank": True}), device=DEV, max_batch=96)
    assert a.raw_bank and not b.raw_bank and a.z.shape[-1] == 2 and b.z.shape[-1] == 3
    ca, cb = a.run(eb, 200, first_frame=77), b.run(eb, 200, first_frame=77)
    assert np.array_equal(ca, cb), (ca, cb)
    assert ca[:, 2].min() == 200 * 2045 and ca[0, 0] > ca[2, 0] > 0


@pytest.mark.parametrize("n_bits", [4096, 1024, 511, 520, 96, 5117, 7000, 10240])
def test_fused_modulate_bank_is_bit_identical_to_the_two_kernel_path(n_bits):
    """wf_cpm_modulate_bank keeps the noisy baseband in shared memory; its bank outputs must equal
    wf_cpm_modulate + wf_mf_bank bit for bit (same samples, same FFMA2 order), at frame lengths that end
    inside a chunk, on a chunk boundary, and in a second scan segment (> 5120 symbols), with per-frame
    noise levels from 0 (noiseless) to 0 dB.  Counters then agree trivially."""
    B = 24
    a = LinkSimulator(LinkConfig(detector="PAM", n_bits=n_bits, extra={"fused": True}), device=DEV, max_batch=B)
    b = LinkSimulator(LinkConfig(detector="PAM", n_bits=n_bits, extra={"fused": False}), device=DEV, max_batch=B)
    assert a.fused and a.rx is None and not b.fused and len(a.stage_names) == 3
    sg = torch.linspace(0.0, 2.0, B, dtype=torch.float32, device=DEV)
    sg[0] = 0.0
    pt = torch.zeros(B, dtype=torch.int32, device=DEV)
    ca = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    cb = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    a.z.fill_(float("nan")); b.z.fill_(float("nan"))
    a.run_batch(12345, sg, pt, ca)
    b.run_batch(12345, sg, pt, cb)
    za = torch.view_as_real(a.z[:, :a.n_calls]).contiguous().view(torch.int32)
    zb = torch.view_as_real(b.z[:, :b.n_calls]).contiguous().view(torch.int32)
    assert a.n_calls == b.n_calls and torch.equal(za, zb), int((za != zb).sum())
    assert torch.equal(ca, cb) and int(ca[0, 2]) == B * a.n_compared


def test_fused_link_covers_soqpsk_tg_and_other_shapes_fall_back_to_four_kernels():
    assert LinkSimulator(LinkConfig(extra={"fused": True}), device=DEV, max_batch=4).fused
    assert LinkSimulator(LinkConfig(), device=DEV, max_batch=4).fused            # the default link since round 2
    assert not LinkSimulator(LinkConfig(extra={"fused": False}), device=DEV, max_batch=4).fused
    for kw in ({"waveform": "soqpsk-mil"}, {"waveform": "soqpsk-a"}, {"detector": "PT"}, {"sps": 10}, {"timing_offset": 3},
               {"waveform": "pcmfm"}, {"extra": {"combine_in_bank": True}}):
        kw = {**kw, "extra": {**kw.get("extra", {}), "fused": True}}
        sim = LinkSimulator(LinkConfig(n_bits=256, **kw), device=DEV, max_batch=4)
        if sim.fused:  # a covered shape must still equal the two-kernel path
            ref = LinkSimulator(LinkConfig(n_bits=256, **{**kw, "extra": {**kw["extra"], "fused": False}}), device=DEV,
                                max_batch=4)
            e = [3.0]
            assert np.array_equal(sim.run(e, 4), ref.run(e, 4)), kw
        else:
            assert "modulate_awgn_bank" not in sim.stage_names and "modulate_awgn" in sim.stage_names, kw
    rc = _ffi.lib().wf_cpm_modulate_bank(None, 1, 16, None, 1, None, 9, 8, 0.0, 0.0, None, None, 0, 0, None, None, 2, 73, 36, 0,
                                         15, None, 16, 1, None)
    assert rc < 0  # bad arguments are refused, not run


def test_link_matches_oracle_chain_with_philox_noise():
    """Same Philox noise through the CUDA chain and through the float64 oracle chain."""
    cfg = LinkConfig(n_bits=1024, detector="PAM")
    sim = LinkSimulator(cfg, device=DEV, max_batch=8)
    sigma = float(o.sigma_from_ebn0_db(4.0, 8))
    sg = torch.full((8,), sigma, dtype=torch.float32, device=DEV)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    sim.run_batch(40, sg, None, counters)
    noise = generate_complex_awgn(sg, sim.n_samples, PhiloxRNG(cfg.seed, first_frame=40), batch=8).cpu().numpy()
    tot = np.zeros(3, dtype=np.int64)
    for b in range(8):
        bits = o.pn_frame_bits(15, 40 + b, 1024)
        tot += np.array(o.soqpsk_link_frame(bits, o.pulse_soqpsk_tg(8), 8, sigma, noise=noise[b].astype(np.complex128),
                                            detector="PAM"))
    got = counters[0].tolist()
    assert got[2] == tot[2]
    # decision level: the same frames stage by stage through the float32 API (same kernels as run_batch),
    # every flip against the float64 oracle gated on the oracle's ACS margin; the fused counters must then
    # equal the errors of exactly these decisions
    from waveforms_b200.glfsr import PNSequence as _PN
    bits = _PN(15).frames(40, 8, 1024)
    syms = TrellisEncoder(tm.SOQPSKTrellis4x2DiffEncoded)(bits)
    rx = wmod.modulate_frames(syms, [0.25], freq_pulse_soqpsk_tg(8), 8, derotate=-np.pi / 4, sigma=sg, seed=cfg.seed,
                              first_frame=40).contiguous()
    db, _ = SOQPSKTrellisDetector().detect(mfbank.pam_bank(rx, freq_pulse_soqpsk_tg(8), 0.25, 8))
    db = db.cpu().numpy()
    errs = 0
    for b in range(8):
        r = o.soqpsk_link_frame(o.pn_frame_bits(15, 40 + b, 1024), o.pulse_soqpsk_tg(8), 8, sigma,
                                noise=noise[b].astype(np.complex128), detector="PAM", return_all=True)
        assert_fp32_flips_are_near_ties(db[b], r["dec_bits"], r["margin"], r["scale"], what=f"Philox frame {40 + b}")
        errs += int(np.count_nonzero(db[b, 2:2 + 1021] != r["bits"][:1021]))
    assert got[0] == errs


def test_link_batching_independence_and_e2e_host_call():
    cfg = LinkConfig(n_bits=512)
    a = LinkSimulator(cfg, device=DEV, max_batch=64).run([3.0, 6.0], 64)
    b = LinkSimulator(cfg, device=DEV, max_batch=10).run([3.0, 6.0], 64)
    assert isinstance(a, np.ndarray) and a.shape == (2, 3) and np.array_equal(a, b)
    assert a[0, 0] > a[1, 0] > 0 and a[0, 2] == 64 * 509
    first = LinkSimulator(cfg, device=DEV, max_batch=64).run([3.0, 6.0], 32)
    second = LinkSimulator(cfg, device=DEV, max_batch=64).run([3.0, 6.0], 32, first_frame=32)
    assert np.array_equal(first + second, a)  # frame sharding (multi-GPU) adds up exactly


DESIGN_EFFECT_FALLBACK = 1.8  # var / binomial of a point's error count when the table carries no per-frame counts


def _reference_design_effect(g, col, k, p_ref):
    """Bit errors of this detector come in bursts (one wrong survivor flips 2-3 neighbouring bits), so the variance
    of a point's error count is LARGER than the binomial n p (1 - p) the Clopper-Pearson interval assumes.  The
    design effect phi = (between-frame variance of the reference's own per-frame error counts) / (binomial variance
    of a frame's count) measures by how much; G10 records the per-frame counts of the replicates generated with
    oracle/make_ber_table.py (errors_by_frame[replicate][point][frame][PT, PAM, PT@-1]).  Floored at 1."""
    if "errors_by_frame" not in g:
        return DESIGN_EFFECT_FALLBACK
    e = g["errors_by_frame"][:, k, :, col].astype(np.float64).ravel()
    n_f = 4093.0  # compared bits per frame (N = 4096, length-2 detector)
    binom = n_f * p_ref * (1.0 - p_ref)
    return max(1.0, float(e.var(ddof=1) / binom)) if binom > 0 and e.size > 1 else DESIGN_EFFECT_FALLBACK


def test_ber_waterfall_inside_reference_confidence_intervals():
    """north_star: simulated BER (Philox noise) inside the REFERENCE's 95 % confidence interval at every Eb/N0
    point.  Three detector rows on the same received frames: PAM and PT at timing offset 0, and PT at offset -1 (the
    operating point the shipped example uses for TG, soqpsk_detection.py:184-185), frames 0..1023 x 13 points,
    noise PCG64(1000 EbN0 + frame + salt).

    The reference table has two parts, pooled:
      * G10  -- the reference ITSELF (oracle/make_ber_table.py), five replicates (salts 0 .. 4e6, ~100 CPU-minutes each);
      * G10b -- REFERENCE-EQUIVALENT replicates (oracle/make_ber_table_oracle.py): the same experiment for further salts
        run through the oracle chain, which reproduces the reference's error count of every frame exactly (CPU test
        test_oracle_link_reproduces_reference_ber_table_frames) and reproduced whole G10 replicates, all 39 cells
        of 4.19e6 bits, before it was trusted (salt 1e6: recorded in G10b as checked_against_g10_salts and in
        profiles/r3_g10b_generation.log; salt 0: a separate run of the same --check).  They are what a longer
        run of the reference would have printed, at 5 minutes a replicate instead of 100.
    Why pool: five replicates are a small sample of the reference's OWN noise -- at 5 dB their mean sits 2.9 of its
    standard errors below the mean of the 20 further replicates (profiles/r2_ber_replicates_4_5dB.txt), and an interval
    centred on that draw rejects the reference's true BER, with which this simulator agrees to 1e-4 relative.  The
    z-scores against G10 alone are printed for the record, not asserted.

    Ours simulates THE SAME 1024 FRAMES 240 times over with independent noise streams (1.0e9 bits per point, ten times
    the pooled table), so the comparison is dominated by the table's sampling error, which is what its interval describes.

    ONE rule, the same in this docstring and in the assert:
      * the interval of a (point, detector) pair is the Clopper-Pearson interval of the pooled counts divided by the
        design effect phi (errors / phi of bits / phi): the standard correction for clustered errors, phi measured
        from the reference's own per-frame counts (_reference_design_effect; 1.1 - 1.8 here).  A plain binomial
        interval would be sqrt(phi) too narrow;
      * the 39 intervals are taken at the Bonferroni level 1 - 0.05/39, i.e. they form a SIMULTANEOUS 95 %
        confidence region for the reference's waterfall, and EVERY pair must fall inside its interval.  (The number
        of pairs inside the plain per-pair 95 % intervals is printed, not asserted: a perfect simulator misses
        about two of 39.)
    A second assert guards against a systematic offset that stays inside every interval: the mean z-score over the
    13 points of a row is N(0, 1/13) for an unbiased simulator; |mean z| < 3.5 / sqrt(13) is required."""
    from waveforms_b200.linksim import ber_confidence_interval

    g = golden("g10_ber_table.npz")
    eb = g["ebn0_db"].astype(float)
    rows = (("PAM", 0, "pam_errors", "compared", 1), ("PT", 0, "pt_errors", "compared", 0),
            ("PT", -1, "pt_m1_errors", "pt_m1_compared", 2))
    n_pairs = len(rows) * eb.size
    level = 1.0 - 0.05 / n_pairs
    n_frames = int(g["frames"])
    n_rep = int(g["replicates"]) if "replicates" in g else 1
    n_seeds = 240  # ten times the bits of the pooled table: our own sampling error adds 5 % to the spread of the difference
    assert n_frames == 1024 and n_pairs == 39
    import os
    from tests.conftest import GOLDEN
    gb = golden("g10b_ber_table_oracle.npz") if os.path.exists(os.path.join(GOLDEN, "g10b_ber_table_oracle.npz")) else None
    n_equiv = 0 if gb is None else int(gb["salts"].size)
    if gb is not None:
        assert str(gb["source"]) == "oracle" and int(gb["frames"]) == n_frames
        assert not set(int(x) for x in gb["salts"]) & set(int(x) for x in g["salts"])       # independent noise
        assert set(int(x) for x in gb["checked_against_g10_salts"]) <= set(int(x) for x in g["salts"])
    print(f"G10: {n_rep} reference replicate(s) + {n_equiv} reference-equivalent (oracle) replicate(s) of {n_frames} frames "
          f"per point; per-frame counts "
          f"{'present' if 'errors_by_frame' in g else 'absent (design effect falls back to %.1f)' % DESIGN_EFFECT_FALLBACK}")
    inside_plain, report = 0, []
    for det, off, key, nkey, col in rows:
        got = np.zeros((eb.size, 3), dtype=np.int64)
        for seed in range(101, 101 + n_seeds):
            sim = LinkSimulator(LinkConfig(detector=det, timing_offset=off, seed=seed), device=DEV, max_batch=13 * 1024)
            got += sim.run(eb, n_frames, first_frame=0)
        assert np.all(got[:, 2] * n_rep == n_seeds * g[nkey])
        zs, zs_ref_only = [], []
        for k in range(eb.size):
            e_own, n_own = int(g[key][k]), int(g[nkey][k])
            e_ref, n_ref = e_own, n_own
            if gb is not None:
                e_ref, n_ref = e_own + int(gb[key][k]), n_own + int(gb[nkey][k])
            p_ref, p_us = e_ref / n_ref, got[k, 0] / got[k, 2]
            phi = _reference_design_effect(g, col, k, p_ref)
            p_own = e_own / n_own
            zs_ref_only.append((p_us - p_own) / np.sqrt(phi * (p_own * (1 - p_own) / n_own + p_us * (1 - p_us) / got[k, 2])))
            lo, hi = ber_confidence_interval(e_ref / phi, n_ref / phi, level)
            lo95, hi95 = ber_confidence_interval(e_ref / phi, n_ref / phi, 0.95)
            inside_plain += lo95 <= p_us <= hi95
            report.append((det, off, eb[k], p_ref, p_us, phi, lo, hi, lo <= p_us <= hi))
            sd = np.sqrt(phi * (p_ref * (1 - p_ref) / n_ref + p_us * (1 - p_us) / got[k, 2]))
            zs.append((p_us - p_ref) / sd)
        print(det, off, "z-scores", np.round(zs, 2).tolist())
        print(det, off, "z-scores against the reference's own replicates alone (not asserted)", np.round(zs_ref_only, 2).tolist())
        zm = float(np.mean(zs))
        assert abs(zm) < 3.5 / np.sqrt(eb.size), (det, off, zm, zs)
    for r in report:
        print("%s off %+d %4.1f dB ref %.4e ours %.4e design effect %.2f simultaneous-95%% CI [%.4e, %.4e] %s" % r)
    print(f"{inside_plain} of {n_pairs} pairs inside the plain (per-pair) 95 % intervals")
    assert all(r[-1] for r in report), [r for r in report if not r[-1]]


# ------------------------------------------------------------------ A11/A12 PCM/FM and multi-h detection
# (no reference detector exists: parity is against the builder's float64 statement oracle/cpm_generic.py)
_GEN = {
    "pcmfm": (lambda: freq_pulse_pcmfm(8, 4), [7], 10, 2, PCMFMSymbolMapper),
    "multih": (lambda: freq_pulse_multih_irig(8), [4, 5], 16, 4, MultiHSymbolMapper),
} if torch.cuda.is_available() else {}


def _cpm_flips_are_near_ties(dec, ref, marg, D, what):
    """CPM detectors: decision k is released at a step < k + 2 D, so it may differ from the float64 statement
    only if some step in [k, k + 2 D) has a float64 margin at float32 resolution."""
    return assert_fp32_flips_are_near_ties(dec, ref, marg[:, 0], marg[:, 1], window=2 * D, what=what)


@pytest.mark.parametrize("name", ["pcmfm", "multih"])
@pytest.mark.parametrize("memory", [1, 2])
def test_generic_cpm_detector_vs_oracle(name, memory, pn15_stream):
    from oracle.cpm_generic import CpmTrellis
    from waveforms_b200.viterbi.cpm import CPMTrellisDetector

    pf, hnum, p, M, mapper = _GEN[name]
    pulse = pf()
    bits = cu(np.stack([pn15_stream[k * 1500:k * 1500 + 1200] for k in range(4)]))
    syms = mapper()(bits)
    bps = 1 if M == 2 else 2
    sg = torch.full((4,), float(o.sigma_from_ebn0_db(5.0, 8, bps)), dtype=torch.float32, device=DEV)
    rx = wmod.modulate_frames(syms, np.array(hnum) / p, pulse, 8, dtype=torch.complex128, derotate=-np.pi / 4,
                              sigma=sg, seed=3, first_frame=0)
    det = CPMTrellisDetector(pulse, hnum, p, M, 8, memory=memory)
    S = syms.shape[1]
    z = det.bank(rx, S)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    dec = det.detect(z, tx_symbols=syms, counters=counters).cpu().numpy()
    tr = CpmTrellis(pulse, hnum, p, M, 8, memory)
    D = tr.default_depth()
    rxh, symh = rx.cpu().numpy(), syms.cpu().numpy()
    errs, refs, margs = 0, [], []
    for b in range(4):
        zref = tr.bank(rxh[b], S)
        assert np.max(np.abs(z[b].cpu().numpy() - zref)) < 1e-9
        r, mg = tr.detect(zref, return_margin=True)
        ref = 2 * r.astype(np.int64) - (M - 1)
        assert np.array_equal(dec[b], ref), (name, memory, b)   # float64 instantiation: identical
        errs += int(np.sum(ref[4:S - 4] != symh[b, 4:S - 4]))
        refs.append(ref)
        margs.append(mg)
    # counters: symbol errors as counted from the decisions, compared column in BITS for every alphabet
    assert counters[0, 1].item() == errs and counters[0, 2].item() == 4 * (S - 8) * bps
    # float32 generic path: flips only at float64 near-ties
    rx32 = rx.to(torch.complex64)
    dec32 = det.detect(det.bank(rx32, S)).cpu().numpy()
    for b in range(4):
        _cpm_flips_are_near_ties(dec32[b], refs[b], margs[b], D, f"{name} L'={memory} generic float32 frame {b}")
    if memory == 2:  # the fused kernel (bank inside the detector, decision bits + traceback), same rule
        assert det.supports_fused(torch.complex64, rx32.shape[1], rx32.shape[1], S)
        cf = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
        decf = det.detect_rx(rx32.contiguous(), S, tx_symbols=syms, counters=cf).cpu().numpy()
        for b in range(4):
            _cpm_flips_are_near_ties(decf[b], refs[b], margs[b], D, f"{name} fused frame {b}")
        se = int(np.sum(decf[:, 4:S - 4] != symh[:, 4:S - 4]))
        assert cf[0, 1].item() == se and cf[0, 2].item() == 4 * (S - 8) * bps
        assert not det.supports_fused(torch.complex128, rx.shape[1], rx.shape[1], S)
    else:
        assert not det.supports_fused(torch.complex64, rx32.shape[1], rx32.shape[1], S)


@pytest.mark.parametrize("name,depth", [("pcmfm", 5), ("pcmfm", 16), ("pcmfm", 32), ("multih", 3), ("multih", 16), ("multih", 32)])
def test_cpm_detector_release_depths(name, depth, pn15_stream):
    """Block-release depth D (survivor memory 2 D): the fused kernels keep decision bits for 2 D <= 64 steps in
    shared memory, the generic kernel 2 D-symbol register-exchange words (2 D log2 M <= 64); both against the
    float64 statement at the same depth, on frame lengths that end inside a block."""
    from oracle.cpm_generic import CpmTrellis
    from waveforms_b200.viterbi.cpm import CPMTrellisDetector

    pf, hnum, p, M, mapper = _GEN[name]
    pulse = pf()
    bps = 1 if M == 2 else 2
    bits = cu(np.stack([pn15_stream[k * 700:k * 700 + 602] for k in range(3)]))
    syms = mapper()(bits)
    sg = torch.full((3,), float(o.sigma_from_ebn0_db(3.0, 8, bps)), dtype=torch.float32, device=DEV)
    rx = wmod.modulate_frames(syms, np.array(hnum) / p, pulse, 8, dtype=torch.complex128, derotate=-np.pi / 4,
                              sigma=sg, seed=11, first_frame=0)
    det = CPMTrellisDetector(pulse, hnum, p, M, 8, memory=2, depth=depth)
    S = syms.shape[1]
    tr = CpmTrellis(pulse, hnum, p, M, 8, 2)
    rx32 = rx.to(torch.complex64).contiguous()
    decf = det.detect_rx(rx32, S).cpu().numpy()
    generic_ok = 2 * depth * bps <= 64
    dec64 = det.detect(det.bank(rx, S)).cpu().numpy() if generic_ok else None
    for b in range(3):
        r, mg = tr.detect(tr.bank(rx[b].cpu().numpy(), S), depth=depth, return_margin=True)
        ref = 2 * r.astype(np.int64) - (M - 1)
        if generic_ok:
            assert np.array_equal(dec64[b], ref), (name, depth, b)
        _cpm_flips_are_near_ties(decf[b], ref, mg, depth, f"{name} fused D={depth} frame {b}")
    if not generic_ok:
        with pytest.raises(ValueError, match="survivor memory"):
            det.detect(det.bank(rx, S))


@pytest.mark.parametrize("name", ["pcmfm", "multih"])
def test_generic_link_round_trip_and_waterfall(name):
    """BASELINE configs[1]/[2] at N = 4096: noiseless frames decode without error; the fused detector link and
    the unfused one (wf_mf_bank + wf_viterbi_cpm) count the same frames with statistically equal results; BER
    falls monotonically with Eb/N0 (the external anchor is test_cpm_ber_against_minimum_distance_bound)."""
    sim = LinkSimulator(LinkConfig(waveform=name), device=DEV, max_batch=1024)
    assert sim.detect_fused and sim.z is None and len(sim.stage_names) == 3
    counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    sim.run_batch(77, torch.zeros(1024, dtype=torch.float32, device=DEV), None, counters)
    be, se, n = counters[0].tolist()
    assert (be, se) == (0, 0) and n == 1024 * (sim.n_sym - 8) * sim.bits_per_symbol == 1024 * sim.n_compared
    got = sim.run([3.0, 6.0, 9.0], 256)
    ber = got[:, 0] / got[:, 2]
    assert ber[0] > ber[1] > ber[2] and ber[2] < 1e-2, ber
    slow = LinkSimulator(LinkConfig(waveform=name, extra={"unfused_detector": True}), device=DEV, max_batch=64)
    assert not slow.detect_fused and len(slow.stage_names) == 4
    a, b = sim.run([4.0, 7.0], 64, first_frame=500), slow.run([4.0, 7.0], 64, first_frame=500)
    assert np.array_equal(a[:, 2], b[:, 2])
    # same noise, same rule, float32 both: only near-tie flips may separate the two error counts
    assert np.all(np.abs(a[:, 0] - b[:, 0]) <= np.maximum(3, 0.01 * b[:, 0])), (a, b)


@pytest.mark.parametrize("name", ["pcmfm", "multih"])
def test_cpm_ber_against_minimum_distance_bound(name):
    """SURVEY 8(c) validation (iii), the only external pin the two reference-less detectors can get: the
    simulated waterfall against the textbook high-SNR law P_b ~ Q(sqrt(d2_min Eb/N0)) (Anderson, Aulin &
    Sundberg, "Digital Phase Modulation", 1986, ch. 3) with d2_min computed from the reference's own pulses
    by oracle/cpm_dmin.py -- 2.58 for IRIG PCM/FM (h = 0.7; binary CPFSK's published 2.43 at h = 0.715 is
    reproduced by the same code) and 1.29(6) for ARTM CPM (Perrins & Rice, IEEE Trans. AES 2007, publish 1.29).
    That law puts BER 1e-3 / 1e-5 at 5.7 / 8.5 dB (PCM/FM; Geoghegan's multi-symbol detector figures for PCM/FM
    are 8.4-8.7 dB at 1e-5) and 8.7 / 11.5 dB (ARTM CPM).  The Eb/N0 at which the simulated BER crosses 1e-3 and
    1e-5 must lie within 1 dB of it -- on the right-hand side (a pulse-truncated detector cannot beat MLSE by
    more than the multiplicity factor: 0.3 dB is allowed for it)."""
    from fractions import Fraction

    from oracle.cpm_dmin import dmin2, ebn0_db_for_ber

    if name == "pcmfm":
        d2 = dmin2(o.pulse_pcmfm(8, 4), [Fraction(7, 10)], 2, 8, span=5)[0]
        grid = np.arange(4.0, 10.01, 0.5)
    else:
        d2 = dmin2(o.pulse_multih(8), [Fraction(4, 16), Fraction(5, 16)], 4, 8, span=5)[0]
        grid = np.arange(7.0, 13.01, 0.5)
    sim = LinkSimulator(LinkConfig(waveform=name), device=DEV, max_batch=16384)
    got = sim.run(grid, 8192, first_frame=31)     # 3.3e7 bits per point
    ber = got[:, 0] / got[:, 2]
    assert np.all(np.diff(ber[got[:, 0] >= 20]) < 0), (ber, got[:, 0])
    lb = np.log10(np.maximum(ber, 1e-12))
    for target in (1e-3, 1e-5):
        k = int(np.flatnonzero(lb < np.log10(target))[0])     # first grid point below the target
        assert got[k, 0] >= 20, (name, target, got[:, 0])       # enough errors on both sides of the crossing
        x = grid[k - 1] + (np.log10(target) - lb[k - 1]) / (lb[k] - lb[k - 1]) * (grid[k] - grid[k - 1])
        want = ebn0_db_for_ber(d2, target)
        print(f"{name}: BER {target:g} at {x:.2f} dB simulated, {want:.2f} dB from d2_min = {d2:.3f}")
        assert -0.3 <= x - want <= 1.0, (name, target, x, want)
