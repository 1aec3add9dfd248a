"""GPU: SURVEY 8(f)3/4 -- the arrays behind the reference's diagnostic figures (waveforms/viz) and the
receive-side filter / sampling-phase estimate (waveforms/filters), CUDA kernels of csrc/diag.cu against
oracle/viz_ref.py and the reference-generated golden G12."""
import numpy as np
import pytest
import torch

from oracle import cpm_link as o
from oracle import viz_ref as vr
from tests.conftest import golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

if torch.cuda.is_available():
    from waveforms_b200 import mfbank
    from waveforms_b200.cpm import modulate as wmod
    from waveforms_b200.cpm.multih import MULTIH_IRIG_DENOM, MULTIH_IRIG_NUMER, freq_pulse_multih_irig
    from waveforms_b200.cpm.pcmfm import freq_pulse_pcmfm
    from waveforms_b200.cpm.soqpsk import freq_pulse_soqpsk_mil, freq_pulse_soqpsk_tg
    from waveforms_b200.cpm.trellis import model as tm
    from waveforms_b200.cpm.trellis.encoder import TrellisEncoder
    from waveforms_b200.filters import estimate_timing_offset, fir_filter, kaiser_fir_lpf
    from waveforms_b200.viz import (constellation, generate_cpm_phase_tree, plot_eye_diagram, plot_phase_tree,
                                    plot_power_spectral_density)


def _tg_signal(n_bits=4096, batch=3):
    bits = np.unpackbits(golden("g1_pn.npz")["pn15"])
    frames = np.stack([bits[k * 5000:k * 5000 + n_bits] for k in range(batch)])
    syms = np.stack([o.fsm_encode(o.TRELLISES["SOQPSKTrellis4x2DiffEncoded"], f) for f in frames])
    t, s = wmod.cpm_modulate(torch.from_numpy(syms).to(DEV), 0.25, freq_pulse_soqpsk_tg(8), 8)
    return t, s.contiguous()


@pytest.mark.parametrize("nfft,bps", [(1024, 1), (256, 2), (4096, 1)])
def test_welch_psd_vs_mlab_restatement(nfft, bps):
    """viz/psd.py:35-40 = matplotlib.mlab.psd (restated in oracle/viz_ref.py; matplotlib itself is absent):
    every bin within 0.02 dB down to 70 dB below the peak (float32 in-place radix-2 FFT in shared memory)."""
    _t, s = _tg_signal()
    f, pxx = plot_power_spectral_density(s, 8, bps, nfft)
    pxx = pxx.cpu().numpy()
    sh = s.cpu().numpy().astype(np.complex128)
    for b in range(s.shape[0]):
        fr, ref = vr.psd(sh[b], 8, bps, nfft)
        assert np.array_equal(f, fr)
        keep = ref > ref.max() * 1e-7
        assert np.max(np.abs(10 * np.log10(pxx[b][keep] / ref[keep]))) < 0.02, (nfft, b)
        assert np.max(np.abs(pxx[b] - ref)) < 1e-5 * ref.max()
    f1, p1 = plot_power_spectral_density(s[0], 8, bps, nfft)   # 1-D input: the reference's call shape
    assert p1.shape == (nfft,) and torch.equal(p1, torch.from_numpy(pxx[0]).to(DEV))
    with pytest.raises(ValueError):
        plot_power_spectral_density(s[0, :100], 8, 1, nfft)
    with pytest.raises(TypeError):
        plot_power_spectral_density(s[0].cpu(), 8, 1, nfft)     # no CPU path


def test_eye_and_constellation_arrays():
    t, s = _tg_signal(1024, 2)
    for sps_mod in ((8, 4), (8, 2), (8, 5)):
        tt, tr = plot_eye_diagram(t, s, *sps_mod)
        for b in range(2):
            rt, ref = vr.eye_traces(t.cpu().numpy(), s[b].cpu().numpy(), *sps_mod)
            assert np.array_equal(tr[b].cpu().numpy(), ref) and np.allclose(tt.cpu().numpy(), rt[0])
    i, q = constellation(s[0], 1000)
    ri, rq = vr.constellation(s[0].cpu().numpy(), 1000)
    assert np.array_equal(i.cpu().numpy(), ri) and np.array_equal(q.cpu().numpy(), rq)


def test_phase_tree_arrays():
    _t, s = _tg_signal(1024, 1)
    sh = s[0].cpu().numpy().astype(np.complex128)
    for modulo, off in ((4, None), (3, 0.5), (13, None)):
        t, ph = plot_phase_tree(s[0], 8, off, modulo)
        rt, ref = vr.phase_tree(sh, 8, off, modulo)
        assert np.array_equal(t, rt) and ph.shape == ref.shape
        assert np.max(np.abs(ph.cpu().numpy() - ref)) < 5e-5   # float32 atan2 on float32 samples


@pytest.mark.parametrize("lab", ["pcmfm", "multih", "mil", "tg"])
def test_generate_cpm_phase_tree_vs_reference(lab):
    """viz/tree.py:89-146 end to end against G12 (the reference's encoder + modulator + unwrap): same symbol
    sequences in the same order, phases within 1e-4 rad (float32 modulator: 1e-5 on the unit circle)."""
    g = golden("g12_phase_tree.npz")
    tre, h, pf = {"pcmfm": (tm.SimpleTrellis2, 0.7, freq_pulse_pcmfm(8, 4)),
                  "multih": (tm.SimpleTrellis4, MULTIH_IRIG_NUMER / MULTIH_IRIG_DENOM, freq_pulse_multih_irig(8)),
                  "mil": (tm.SOQPSKTrellis4x2DiffEncoded, 0.25, freq_pulse_soqpsk_mil(8)),
                  "tg": (tm.SOQPSKTrellis4x2DiffEncoded, 0.25, freq_pulse_soqpsk_tg(8))}[lab]
    t, ph = generate_cpm_phase_tree(pf, h, TrellisEncoder(tre), 8)
    ref = g[f"{lab}_phases"]
    assert ph.shape == ref.shape and t.size == ref.shape[1]
    assert np.max(np.abs(ph.cpu().numpy() - ref)) < 1e-4


def test_kaiser_lpf_front_end_and_timing_estimate():
    """filters/lpf.py:13-34 design (identical taps), its batched application (np.convolve 'same' per frame), and
    the sampling-phase estimate that replaces the hard-coded offsets of soqpsk_detection.py:183-192: on a
    noiseless SOQPSK-TG frame delayed by d samples the estimate moves by d, the peak of the undelayed frame is within
    one sample of the reference's offset 0 with the reference's 0 / -1 on its shoulder, and noise at the link's
    operating point does not move it."""
    taps = kaiser_fir_lpf(8, 1.0)
    assert np.array_equal(taps, vr.kaiser_fir_lpf(8, 1.0))
    _t, s = _tg_signal(2048, 2)
    rng = np.random.default_rng(0)
    noisy = s + torch.from_numpy((0.3 * (rng.normal(size=tuple(s.shape)) + 1j * rng.normal(size=tuple(s.shape)))).astype(np.complex64)).to(DEV)
    y = fir_filter(noisy, taps)
    nh = noisy.cpu().numpy().astype(np.complex128)
    for b in range(2):
        assert np.max(np.abs(y[b].cpu().numpy() - np.convolve(nh[b], taps, "same"))) < 2e-5
    rho0 = np.asarray(mfbank.pam_taps(freq_pulse_soqpsk_tg(8), 0.25, 8)[0][0])
    rx = (s * complex(np.exp(-1j * np.pi / 4))).contiguous()
    base = None
    for d in (0, 1, 3, 6):
        delayed = torch.roll(rx, d, dims=1)
        mf = fir_filter(delayed, rho0)
        off, met = estimate_timing_offset(mf, 8)
        assert np.allclose(met.cpu().numpy(), np.stack([vr.timing_metric(m, 8) for m in mf.cpu().numpy().astype(np.complex128)]),
                           rtol=1e-4, atol=1e-6)
        ph = (-off) % 8
        if base is None:
            base = ph
            # the peak sits within one sample of the reference's offset 0 (the 'same' convolution puts it just
            # before n = 0), and the reference's hard-coded 0 and -1 are on its shoulder, well above the trough
            assert torch.all((off == 0) | (off == 1)), off
            m = met.cpu().numpy()
            assert np.all(m[:, 0] > 0.85 * m.max(axis=1)) and np.all(m[:, 1] > 0.6 * m.max(axis=1)), m
            assert np.all(m.min(axis=1) < 0.25 * m.max(axis=1)), m
        assert torch.all(ph == (base + d) % 8), (d, ph, base)
    # the estimate holds at Eb/N0 = 6 dB (sigma 1.0 per component at sps 8 ... ~0 dB per sample)
    noisy_rx = rx + torch.from_numpy((1.0 * (rng.normal(size=tuple(rx.shape)) + 1j * rng.normal(size=tuple(rx.shape)))).astype(np.complex64)).to(DEV)
    off_n, _ = estimate_timing_offset(fir_filter(noisy_rx.contiguous(), rho0), 8)
    assert torch.all((-off_n) % 8 == base), (off_n, base)
