"""Dibit -> {-3,-1,+1,+3} mapper on the GPU -- mirrors waveforms/cpm/multih/precoder.py:5-23
(first bit of each pair is the MSB; odd-length input is a ValueError)."""
from __future__ import annotations

import torch

from ... import _dev, _ffi


class MultiHSymbolMapper:
    def __init__(self) -> None:
        self.i = 0

    def __call__(self, bits: torch.Tensor) -> torch.Tensor:
        bits = _dev.require_cuda(bits, "bits", torch.uint8)
        frames, was_1d = _dev.as_batch(bits)
        B, N = frames.shape
        if N % 2:
            msg = "Odd length bit array passed into quaternary mapper."
            raise ValueError(msg)
        syms = torch.empty((B, N // 2), dtype=torch.int8, device=frames.device)
        with torch.cuda.device(frames.device):
            _ffi.check(_ffi.lib().wf_bits_encode(_dev.ptr(frames), B, N, _ffi.ENC_QUATERNARY, None, None, 0,
                                                 _dev.ptr(syms), _dev.stream_ptr(frames.device)))
        self.i = (self.i + N) % 2
        return syms[0] if was_1d else syms
