"""CPU: host-side logic of the product (pulse / tap design, trellis tables, LFSR algebra
in the C library, argument checking, sharding arithmetic) and the C-ABI surface.  No
kernel is launched here; compute calls must fail loudly without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import cpm_link as o
from tests.conftest import ROOT, golden
from waveforms_b200 import _ffi, linksim, mfbank
from waveforms_b200.cpm.multih import MULTIH_IRIG_DENOM, MULTIH_IRIG_NUMER, freq_pulse_multih_irig
from waveforms_b200.cpm.pamapprox import pam_unit_pulse2, rho_pulses
from waveforms_b200.cpm.pcmfm import PCMFM_DENOM, PCMFM_NUMER, freq_pulse_pcmfm
from waveforms_b200.cpm.soqpsk import (SOQPSK_DENOM, SOQPSK_NUMER, freq_pulse_soqpsk_a, freq_pulse_soqpsk_b,
                                      freq_pulse_soqpsk_mil, freq_pulse_soqpsk_tg)
from waveforms_b200.cpm.trellis import model as tm
from waveforms_b200.cpm.trellis.encoder import TrellisEncoder
from waveforms_b200.glfsr.pn import GALOIS_LFSR_POLYS, PNSequence, generate_mask


# ---- C ABI surface -------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "wfb200.h")).read()
    declared = set(re.findall(r"WF_API\s+[\w\s\*]+?\b(wf_\w+)\s*\(", hdr))
    assert len(declared) >= 16
    lib = ctypes.CDLL(_ffi.lib_path()) if os.path.exists(_ffi.lib_path()) else _ffi.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in wfb200.h but not exported"
    assert declared == set(_ffi.PROTOTYPES), declared ^ set(_ffi.PROTOTYPES)
    assert _ffi.lib().wf_abi_version() == _ffi.ABI_VERSION


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    lib = _ffi.lib()
    assert lib.wf_device_check() == -2
    assert b"no CPU fallback" in lib.wf_last_error_string()
    with pytest.raises(TypeError, match="no CPU path"):
        TrellisEncoder(tm.SimpleTrellis2)(torch.zeros(8, dtype=torch.uint8))
    with pytest.raises(RuntimeError):
        PNSequence(9).generate_sequence()
    syms = np.zeros(4, np.int8)
    pulse = (ctypes.c_double * 3)(0.5, 1.0, 0.5)
    h = (ctypes.c_double * 1)(0.5)
    out = np.zeros(40, np.complex64)
    rc = lib.wf_cpm_modulate(syms.ctypes.data, 1, 4, h, 1, pulse, 3, 8, 0.0, 0.0, 0, out.ctypes.data, 40, None, None, 0,
                             0, None, None)
    assert rc == -2 and not out.any()


def test_argument_errors_are_reported():
    lib = _ffi.lib()
    rc = lib.wf_prbs_encode(0x6000, 15, 0x7FFF, 0, None, 0, 1, 33, _ffi.ENC_QUATERNARY, None, None, None, 0, None, None)
    assert rc == -1 and b"Odd length" in lib.wf_last_error_string()
    rc = lib.wf_prbs_encode(0x6000, 70, 0x7FFF, 0, None, 0, 1, 32, 0, None, None, None, 0, None, None)
    assert rc == -1 and b"degree" in lib.wf_last_error_string()
    with pytest.raises(ValueError):
        _ffi.check(rc)


# ---- A1 host algebra ----------------------------------------------------------------
def test_masks_and_keyerror():
    g = golden("g1_pn.npz")
    assert np.array_equal(np.array([generate_mask(d) for d in range(2, 65)], dtype=np.uint64), g["masks"])
    assert GALOIS_LFSR_POLYS[15] == [15, 14] and GALOIS_LFSR_POLYS[8] == [8, 6, 5, 4] and len(GALOIS_LFSR_POLYS) == 65
    for bad in (0, 1, 65, -3):
        with pytest.raises(KeyError, match="PRBS Polynomial Not Defined"):
            generate_mask(bad)


@pytest.mark.parametrize("deg", [2, 9, 15, 23, 33, 64])
def test_lfsr_jump_ahead(deg):
    lib = _ffi.lib()
    mask, st0 = o.pn_mask(deg), (1 << deg) - 1
    for k in (0, 1, 31, 32, 1000, 12345):
        _, st = o.lfsr_bits(mask, st0, k)
        assert lib.wf_lfsr_jump(mask, deg, st0, k) == st
    if deg <= 23:  # x^(2^n - 1) = 1: a full period returns to the start
        assert lib.wf_lfsr_jump(mask, deg, st0, (1 << deg) - 1) == st0
    a = lib.wf_lfsr_jump(mask, deg, st0, 10 ** 15)
    assert lib.wf_lfsr_jump(mask, deg, lib.wf_lfsr_jump(mask, deg, st0, 10 ** 15 - 77), 77) == a


# ---- A2 trellis tables, A5/A9 pulse design ---------------------------------------------------
def test_trellis_tables_match_reference_listing():
    g = golden("g9_trellis.npz")
    for name in ("SOQPSKTrellis8x1", "SOQPSKTrellis4x2", "SOQPSKTrellis4x2DiffEncoded", "SimpleTrellis2", "SimpleTrellis4"):
        t = getattr(tm, name)
        flat = np.array([[(b.inp, b.out, b.start, b.end) for b in col] for col in t.branches])
        assert np.array_equal(flat, g[name]), name  # including branch ORDER (tie-break semantics)
        assert (t.columns, t.states, t.branches_per_column, t.input_cardinality, t.output_cardinality) == tuple(g[name + "_shape"])
        cols = [[(b.inp, b.out, b.start, b.end) for b in col] for col in t.branches]
        assert cols == [[tuple(b) for b in c] for c in o.TRELLISES[name]]
    fsm = tm.FiniteStateMachine(tm.SOQPSKTrellis4x2DiffEncoded)
    assert fsm.symbols == [-2, 0, 2] and fsm.symbol_idx_map == {-2: 0, 0: 1, 2: 2}
    assert fsm.reverse_transitions[0][0][2].inp == 1 and fsm.forward_branch_mapping[1][1][1].out == 2
    with pytest.raises(ValueError, match="multiple of 2"):
        tm.Trellis(branches=[[tm.Branch(0, 0, 0, 0), tm.Branch(1, 0, 0, 0), tm.Branch(2, 0, 0, 0)]]).input_cardinality


@pytest.mark.parametrize("sps", [8, 10])
def test_pulses_and_rho(sps):
    g = golden("g3_pulses.npz")
    cases = {"tg": freq_pulse_soqpsk_tg(sps), "a": freq_pulse_soqpsk_a(sps), "b": freq_pulse_soqpsk_b(sps),
             "mil": freq_pulse_soqpsk_mil(sps), "pcmfm4": freq_pulse_pcmfm(sps, 4), "pcmfm6": freq_pulse_pcmfm(sps, 6),
             "multih": freq_pulse_multih_irig(sps)}
    for k, v in cases.items():
        np.testing.assert_allclose(v, g[f"{k}_{sps}"], rtol=0, atol=1e-14, err_msg=k)
    for lab, f in (("tg", freq_pulse_soqpsk_tg), ("mil", freq_pulse_soqpsk_mil)):
        r = rho_pulses(f(sps), 0.25, sps, 2)
        np.testing.assert_allclose(r[0], g[f"rho0_{lab}_{sps}"], rtol=0, atol=1e-14)
        np.testing.assert_allclose(r[1], g[f"rho1_{lab}_{sps}"], rtol=0, atol=1e-14)
    assert (SOQPSK_NUMER, SOQPSK_DENOM, PCMFM_NUMER, PCMFM_DENOM, MULTIH_IRIG_DENOM) == (1, 4, 7, 10, 16)
    assert list(MULTIH_IRIG_NUMER) == [4, 5]
    q = np.cumsum(freq_pulse_soqpsk_tg(sps)) / sps
    assert pam_unit_pulse2(q, 0.25).size == 2 * q.size - 1


def test_bank_tables_and_instants():
    pulse = freq_pulse_soqpsk_tg(8)
    np.testing.assert_allclose(mfbank.pt_taps(pulse, 0.25, 8), o.pt_taps(pulse, 0.25, 8), rtol=0, atol=1e-15)
    taps, comb = mfbank.pam_taps(pulse, 0.25, 8)
    assert taps.shape == (2, 73) and comb.shape == (3, 2) and not taps[1, 65:].any()
    np.testing.assert_allclose(comb, np.conj(o.PSEUDO_SYMBOLS).T, rtol=0, atol=0)
    for n, sps, off, length in ((32776, 8, 0, 2), (327690, 10, -1, 2), (327690, 10, -3, 2), (4104, 8, 0, 4), (100, 8, 3, 2),
                                (16, 8, 0, 2), (17, 8, 0, 2)):
        inst = o.symbol_instants(n, sps, off, length)
        first, calls = mfbank.symbol_instants(n, sps, off, length)
        assert calls == inst.size and (calls == 0 or first == inst[0]), (n, sps, off, length)


def test_fused_modulate_bank_shape_query():
    """wf_cpm_modulate_bank_supported is pure host arithmetic: the SOQPSK-TG / PAM shape at sps 8 is covered for
    every frame length, anything else is refused (the caller then runs wf_cpm_modulate + wf_mf_bank); the compute
    entry point itself fails loudly without a GPU."""
    lib = _ffi.lib()
    K = freq_pulse_soqpsk_tg(8).size
    for S in (1, 95, 511, 512, 4096, 5119, 5120, 7000, 100000):
        first, calls = mfbank.symbol_instants((S + 1) * 8, 8, 0, 2)
        assert lib.wf_cpm_modulate_bank_supported(K, 8, 2, 73, 36, first, S, calls) == 1, S
    first, calls = mfbank.symbol_instants(4097 * 8, 8, 0, 2)
    assert lib.wf_cpm_modulate_bank_supported(K, 8, 2, 73, 36, first, 4096, 4600) == 1      # last call the chunks cover
    assert lib.wf_cpm_modulate_bank_supported(K, 8, 2, 73, 36, first, 4096, 4601) == 0      # one call too many
    assert lib.wf_cpm_modulate_bank_supported(K, 10, 2, 91, 45, first, 4096, calls) == 0     # sps 10
    assert lib.wf_cpm_modulate_bank_supported(K, 8, 3, 9, 4, first, 4096, calls) == 0        # PT bank
    assert lib.wf_cpm_modulate_bank_supported(K, 8, 2, 73, 36, 5, 4096, calls) == 0          # odd tile origin
    assert lib.wf_cpm_modulate_bank_supported(freq_pulse_pcmfm(8).size, 8, 2, 73, 36, first, 4096, calls) == 0
    assert lib.wf_cpm_modulate_bank_supported(freq_pulse_soqpsk_b(8).size, 8, 2, 73, 36, first, 4096, calls) == 0
    if not torch.cuda.is_available():
        h, g, t = _ffi.dbl_array([0.25]), _ffi.dbl_array(freq_pulse_soqpsk_tg(8)), _ffi.dbl_array(np.zeros(146))
        rc = lib.wf_cpm_modulate_bank(ctypes.c_void_p(16), 1, 16, h, 1, g, K, 8, 0.0, 0.0, None, None, 0, 0, None, t, 2, 73, 36, 0,
                                      15, ctypes.c_void_p(16), 16, 1, None)
        assert rc == -2  # WF_ERR_DEVICE



def test_link_cfg_struct_and_workspace_query():
    """struct wf_link_cfg is part of the ABI: the ctypes mirror has the C layout (static_assert in csrc/link.cu says
    152 bytes), and wf_link_sim_workspace_bytes is pure host arithmetic (packed bits | symbols | raw bank outputs,
    each rounded up to 256 bytes); bad configurations come back as negative wf_status with a message."""
    lib = _ffi.lib()
    assert ctypes.sizeof(_ffi.LinkCfg) == 152
    one = np.ones(6)
    cfg = _ffi.LinkCfg(mask=0x6000, init_state=32767, period=32767, degree=15, encoder=_ffi.ENC_SOQPSK_DIFF, n_bits=4096,
                       n_filt=2, n_taps=73, center=36, n_calls=4097, sps=8, combine_re=one.ctypes.data,
                       combine_im=one.ctypes.data)
    up = lambda v: (v + 255) // 256 * 256  # noqa: E731
    for B in (0, 1, 7, 4096):
        want = up(B * 128 * 4) + up(B * 4096) + up(B * 4100 * 16)
        assert lib.wf_link_sim_workspace_bytes(ctypes.byref(cfg), B) == want, B
    cfg.encoder = _ffi.ENC_QUATERNARY
    assert lib.wf_link_sim_workspace_bytes(ctypes.byref(cfg), 4) == -1 and b"SOQPSK" in lib.wf_last_error_string()
    cfg.encoder, cfg.n_filt = _ffi.ENC_SOQPSK, 3
    assert lib.wf_link_sim_workspace_bytes(ctypes.byref(cfg), 4) == -1 and b"PAM bank" in lib.wf_last_error_string()
    assert lib.wf_link_sim_workspace_bytes(None, 4) == -1
    assert lib.wf_static_scratch_max_calls() == (32 << 20) // 24


# ---- sweep arithmetic ---------------------------------------------------------------------
def test_shard_range_partitions():
    for n in (0, 1, 7, 8, 65536, 1000003):
        for world in (1, 2, 3, 8):
            spans = [linksim.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


def test_confidence_interval_and_sigma():
    lo, hi = linksim.ber_confidence_interval(7, 392928)
    assert 7.0e-6 < lo < 7.3e-6 and 3.6e-5 < hi < 3.7e-5
    assert linksim.ber_confidence_interval(0, 1000)[0] == 0.0
    from waveforms_b200.noise import sigma_from_ebn0_db
    assert abs(sigma_from_ebn0_db(6.0, 8) - 1.0024) < 1e-4
    assert abs(sigma_from_ebn0_db(10.0, 10) - np.sqrt(2) / 2) < 1e-12  # the example's sigma
    assert np.allclose(sigma_from_ebn0_db([0, 3], 8, 2), o.sigma_from_ebn0_db([0, 3], 8, 2))
