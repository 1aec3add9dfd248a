"""GPU: wf_soqpsk_detect_static (csrc/noalloc.cu), the allocation-free variant of bank + SOQPSK detector whose
intermediate lives in a static device array inside the library, against the existing two-call form
(wf_mf_bank + wf_viterbi_soqpsk with a caller buffer), through the C ABI.  Integer outputs, so the
tolerance is the one every detector test here uses: bit-exact decisions and counters.  Two calls in a
row with different shapes (the second smaller, so the array still holds the first call's values where
the second does not write), plus the size limit.  Collected last on purpose;
it passes alone in a fresh process and after the other tests."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _two_call(sim, lib, _dev, _ffi, B, counters, dec_bits, dec_syms, z):
    cfg = sim.cfg
    st = _dev.stream_ptr(sim.device)
    nt = sim._taps.shape[1]
    _ffi.check(lib.wf_mf_bank(_dev.ptr(sim.rx), _ffi.WF_F32, B, sim.n_samples, sim.ld_rx, _dev.f64_ptr(sim._tap_re),
                              None if sim._tap_im is None else _dev.f64_ptr(sim._tap_im), sim._taps.shape[0], nt,
                              (nt - 1) // 2, None if sim._comb_re is None else _dev.f64_ptr(sim._comb_re),
                              None if sim._comb_im is None else _dev.f64_ptr(sim._comb_im), 3, cfg.sps,
                              sim.first_instant, sim.n_calls, _dev.ptr(z), sim.n_calls, st))
    _ffi.check(lib.wf_viterbi_soqpsk(_dev.ptr(z), _ffi.WF_F32, B, sim.n_calls, sim.n_calls, int(cfg.differential),
                                     cfg.length, _dev.ptr(dec_bits), _dev.ptr(dec_syms), _dev.ptr(sim.packed),
                                     sim.ld_words, _dev.ptr(sim.syms), sim.n_sym, None, _dev.ptr(counters), None, st))


def _static_call(sim, lib, _dev, _ffi, B, counters, dec_bits, dec_syms):
    cfg = sim.cfg
    st = _dev.stream_ptr(sim.device)
    nt = sim._taps.shape[1]
    return lib.wf_soqpsk_detect_static(
        _dev.ptr(sim.rx), B, sim.n_samples, sim.ld_rx, _dev.f64_ptr(sim._tap_re),
        None if sim._tap_im is None else _dev.f64_ptr(sim._tap_im), sim._taps.shape[0], nt, (nt - 1) // 2,
        None if sim._comb_re is None else _dev.f64_ptr(sim._comb_re),
        None if sim._comb_im is None else _dev.f64_ptr(sim._comb_im), cfg.sps, sim.first_instant, sim.n_calls,
        int(cfg.differential), cfg.length, _dev.ptr(dec_bits), _dev.ptr(dec_syms), _dev.ptr(sim.packed), sim.ld_words,
        _dev.ptr(sim.syms), sim.n_sym, None, _dev.ptr(counters), st)


@pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")
def test_static_scratch_detector_equals_two_call_form_twice_in_a_row():
    os.environ["WF_STATIC_SCRATCH_POISON"] = "1"  # read once, at the entry point's first call in this process
    from waveforms_b200 import _dev, _ffi
    from waveforms_b200.linksim import LinkConfig, LinkSimulator

    lib = _ffi.lib()
    limit = int(lib.wf_static_scratch_max_calls())
    assert limit == (32 << 20) // 24
    # (detector, frame bits, frames, Eb/N0): the second call is SMALLER than the first, and a different bank
    for det, n_bits, B, ebn0 in (("PAM", 2048, 96, 3.0), ("PT", 512, 40, 1.0), ("PAM", 2048, 96, 3.0)):
        sim = LinkSimulator(LinkConfig(n_bits=n_bits, detector=det, seed=77, extra={"fused": False, "combine_in_bank": True}),
                            device=DEV, max_batch=B)
        assert sim.rx is not None and B * sim.n_calls <= limit
        got = sim.run([ebn0], B, first_frame=5)  # leaves the B noisy frames in sim.rx, their bits in sim.packed
        assert got[0, 0] > 0  # errors present: the comparison below is not vacuous
        z = torch.empty((B, sim.n_calls, 3), dtype=torch.complex64, device=DEV)
        want_c = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
        want_b = torch.zeros((B, sim.n_calls), dtype=torch.uint8, device=DEV)
        want_s = torch.zeros((B, sim.n_calls), dtype=torch.int8, device=DEV)
        _two_call(sim, lib, _dev, _ffi, B, want_c, want_b, want_s, z)
        have_c, have_b, have_s = torch.zeros_like(want_c), torch.zeros_like(want_b), torch.zeros_like(want_s)
        before = torch.cuda.memory_allocated(DEV)
        _ffi.check(_static_call(sim, lib, _dev, _ffi, B, have_c, have_b, have_s))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(DEV) == before
        assert np.array_equal(have_c.cpu().numpy(), want_c.cpu().numpy()), (det, have_c, want_c)
        # LinkSimulator.run counted the same frames (it does not ask for the symbol-error column)
        assert np.array_equal(want_c.cpu().numpy()[:, [0, 2]], got[:, [0, 2]]) and int(want_c[0, 1]) > 0, (det, want_c, got)
        assert torch.equal(have_b, want_b) and torch.equal(have_s, want_s)
    # beyond the limit: refused with a message, nothing launched, counters untouched
    sim = LinkSimulator(LinkConfig(n_bits=512, detector="PAM", extra={"fused": False, "combine_in_bank": True}),
                        device=DEV, max_batch=8)
    c = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    too_many = limit // sim.n_calls + 1
    rc = _static_call(sim, lib, _dev, _ffi, too_many, c, torch.zeros(1, dtype=torch.uint8, device=DEV),
                      torch.zeros(1, dtype=torch.int8, device=DEV))
    assert rc == -1 and b"static scratch" in lib.wf_last_error_string()
    with pytest.raises(ValueError, match="static scratch"):
        _ffi.check(rc)
    torch.cuda.synchronize()
    assert int(c.abs().sum()) == 0
