"""Closed-form ternary SOQPSK precoder on the GPU -- mirrors
waveforms/cpm/soqpsk/precoder.py:5-24: alpha_i = (-1)^(i+1) (2 b_{i-1} - 1)(b_i - b_{i-2})
with outputs in {-1, 0, +1}, two bits of memory and the parity ``i`` carried
across calls.  Runs through wf_bits_encode (WF_ENC_SOQPSK_TERNARY)."""
from __future__ import annotations

import torch

from ... import _dev, _ffi


class SOQPSKPrecoder:
    def __init__(self) -> None:
        self.i = 0
        self.mem = 0, 0

    def __call__(self, bits: torch.Tensor) -> torch.Tensor:
        bits = _dev.require_cuda(bits, "bits", torch.uint8)
        frames, was_1d = _dev.as_batch(bits)
        B, N = frames.shape
        syms = torch.empty((B, N), dtype=torch.int8, device=frames.device)
        # kernel state = (index parity, 2*latest even-indexed bit + latest odd-indexed bit)
        m2, m1 = int(self.mem[0]), int(self.mem[1])  # b_{i-2}, b_{i-1}
        st = (2 * m2 + m1) if self.i % 2 == 0 else (2 * m1 + m2)
        state = torch.tensor([[self.i, st]], dtype=torch.int32, device=frames.device).repeat(B, 1)
        with torch.cuda.device(frames.device):
            _ffi.check(_ffi.lib().wf_bits_encode(_dev.ptr(frames), B, N, _ffi.ENC_SOQPSK_TERNARY, _dev.ptr(state), None,
                                                 0, _dev.ptr(syms), _dev.stream_ptr(frames.device)))
        if was_1d:
            tail = ([m2, m1] + frames[0, -2:].tolist())[-2:]
            self.mem = tail[0], tail[1]
            self.i = (self.i + N) % 2
            return syms[0]
        return syms
