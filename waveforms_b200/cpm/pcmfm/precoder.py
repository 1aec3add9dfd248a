"""PCM/FM bit -> {-1,+1} mapper on the GPU -- mirrors waveforms/cpm/pcmfm/precoder.py:5-15."""
from __future__ import annotations

import torch

from ... import _dev, _ffi


class PCMFMSymbolMapper:
    def __call__(self, bits: torch.Tensor) -> torch.Tensor:
        bits = _dev.require_cuda(bits, "bits", torch.uint8)
        frames, was_1d = _dev.as_batch(bits)
        B, N = frames.shape
        syms = torch.empty((B, N), dtype=torch.int8, device=frames.device)
        with torch.cuda.device(frames.device):
            _ffi.check(_ffi.lib().wf_bits_encode(_dev.ptr(frames), B, N, _ffi.ENC_BINARY, None, None, 0, _dev.ptr(syms),
                                                 _dev.stream_ptr(frames.device)))
        return syms[0] if was_1d else syms
