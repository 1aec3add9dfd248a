"""GPU: wf_link_sim_fused (csrc/link.cu), the whole SOQPSK link in one C call with a caller-provided workspace,
against the three separate calls it stands for (wf_prbs_encode -> wf_cpm_modulate_bank ->
wf_viterbi_soqpsk_combined).  Integer counters: bit-exact."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")
@pytest.mark.parametrize("wf,n_bits,diff", [("soqpsk-tg", 1024, True), ("soqpsk-tg", 4096, False), ("soqpsk-a", 520, True)])
def test_one_call_link_equals_three_calls(wf, n_bits, diff):
    from waveforms_b200.linksim import LinkConfig, LinkSimulator

    eb = [1.0, 4.0, 7.0]
    one = LinkSimulator(LinkConfig(waveform=wf, n_bits=n_bits, differential=diff, seed=9), device=DEV, max_batch=48)
    three = LinkSimulator(LinkConfig(waveform=wf, n_bits=n_bits, differential=diff, seed=9, extra={"one_call": False}),
                          device=DEV, max_batch=48)
    assert one.fused and one._ws is not None and one.one_call and not three.one_call
    a = one.run(eb, 130, first_frame=77)     # 390 rows in batches of 48: ragged last batch, mixed points per batch
    b = three.run(eb, 130, first_frame=77)
    assert np.array_equal(a, b), (a, b)
    assert np.all(a[:, 2] == 130 * one.n_compared)
    if diff:
        assert a[0, 0] > a[2, 0] > 0
    else:
        # the non-differential trellis in this chain reads BER 1/2 in the reference too (its noiseless replay of
        # soqpsk_detection.py with SOQPSKTrellis4x2 gives 511 errors of 1021): parity, not a defect of the link
        assert np.all(np.abs(a[:, 0] / a[:, 2] - 0.5) < 0.02)
    # the un-fused four-kernel path counts the same frames identically too
    four = LinkSimulator(LinkConfig(waveform=wf, n_bits=n_bits, differential=diff, seed=9, extra={"fused": False}),
                         device=DEV, max_batch=64)
    assert np.array_equal(four.run(eb, 130, first_frame=77), a)


@pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")
def test_one_call_link_argument_checks():
    from waveforms_b200 import _dev, _ffi
    from waveforms_b200.linksim import LinkConfig, LinkSimulator

    lib = _ffi.lib()
    sim = LinkSimulator(LinkConfig(n_bits=512), device=DEV, max_batch=8)
    need = int(lib.wf_link_sim_workspace_bytes(C.byref(sim._link_cfg), 8))
    assert need == sim._ws.numel() == sim.workspace_bytes()
    sigma = torch.full((8,), 0.5, dtype=torch.float32, device=DEV)
    counters = torch.zeros((1, 3), dtype=torch.int64, device=DEV)
    st = _dev.stream_ptr(sim.device)
    with pytest.raises(ValueError, match="workspace"):
        _ffi.check(lib.wf_link_sim_fused(C.byref(sim._link_cfg), 8, 0, None, None, _dev.ptr(sigma), None, _dev.ptr(sim._ws),
                                         need - 1, _dev.ptr(counters), st))
    bad = _ffi.LinkCfg.from_buffer_copy(sim._link_cfg)
    bad.encoder = _ffi.ENC_BINARY
    with pytest.raises(ValueError, match="SOQPSK"):
        _ffi.check(lib.wf_link_sim_fused(C.byref(bad), 8, 0, None, None, _dev.ptr(sigma), None, _dev.ptr(sim._ws), need,
                                         _dev.ptr(counters), st))
    odd = _ffi.LinkCfg.from_buffer_copy(sim._link_cfg)
    odd.sps = 4  # not a shape the fused kernel is built for
    rc = lib.wf_link_sim_fused(C.byref(odd), 8, 0, None, None, _dev.ptr(sigma), None, _dev.ptr(sim._ws), need,
                               _dev.ptr(counters), st)
    assert rc == -4 and b"not covered" in lib.wf_last_error_string()
    torch.cuda.synchronize()
    assert int(counters.abs().sum()) == 0
    _ffi.check(lib.wf_link_sim_fused(C.byref(sim._link_cfg), 8, 0, None, None, _dev.ptr(sigma), None, _dev.ptr(sim._ws), need,
                                     _dev.ptr(counters), st))
    torch.cuda.synchronize()
    assert int(counters[0, 2]) == 8 * sim.n_compared
