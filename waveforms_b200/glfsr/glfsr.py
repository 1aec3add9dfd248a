"""Galois LFSR bit source on the GPU -- mirrors waveforms/glfsr/glfsr.py:1-19.

``GLFSR(mask, state).next_bit()`` keeps the reference's one-bit-per-call interface;
the bits themselves come from the jump-ahead PRBS kernel (wf_prbs_encode) in blocks
and are handed out from a small host buffer.  ``next_bits(n)`` and ``frames(...)`` are
the batched forms that stay on the device.
"""
from __future__ import annotations

import torch

from .. import _dev, _ffi

_BLOCK = 4096  # bits fetched per kernel call by next_bit()


class GLFSR:
    def __init__(self, mask: int, state: int) -> None:
        self.mask = int(mask)
        self._state = int(state)
        self._buf: list[int] = []

    # the reference exposes .state as a plain attribute (glfsr.py:4); keep it coherent
    # with the buffered bits: state = register contents after the bits handed out so far.
    @property
    def state(self) -> int:
        return self._state

    @state.setter
    def state(self, value: int) -> None:
        self._state = int(value)
        self._buf = []

    @property
    def _degree(self) -> int:
        return max(self.mask.bit_length(), self._state.bit_length(), 1)

    def _check(self) -> int:
        deg = self._degree
        if deg > 64:
            raise ValueError("GLFSR registers wider than 64 bits are not supported")
        return deg

    def next_bits(self, count: int, device=None) -> torch.Tensor:
        """``count`` successive outputs of next_bit() as a uint8 CUDA tensor; advances the state."""
        deg = self._check()
        self._buf = []
        dev = torch.device(device) if device is not None else _dev.default_device()
        out = torch.empty((1, count), dtype=torch.uint8, device=dev)
        if count:
            with torch.cuda.device(dev):
                _ffi.check(_ffi.lib().wf_prbs_encode(self.mask, deg, self._state, 0, None, 0, 1, count,
                                                     _ffi.ENC_BINARY, None, _dev.ptr(out), None, 0, None,
                                                     _dev.stream_ptr(dev)))
            self._state = int(_ffi.lib().wf_lfsr_jump(self.mask, deg, self._state, count))
        return out[0]

    def next_bit(self) -> int:
        """Produces the next bit in the GLFSR sequence (glfsr.py:6-19)."""
        if not self._buf:
            deg = self._check()
            dev = _dev.default_device()
            blk = torch.empty((1, _BLOCK), dtype=torch.uint8, device=dev)
            _ffi.check(_ffi.lib().wf_prbs_encode(self.mask, deg, self._state, 0, None, 0, 1, _BLOCK,
                                                 _ffi.ENC_BINARY, None, _dev.ptr(blk), None, 0, None,
                                                 _dev.stream_ptr(dev)))
            self._buf = blk[0].cpu().tolist()[::-1]
        bit = self._buf.pop()
        # one register step on the host keeps .state exact between buffered bits
        self._state = int(_ffi.lib().wf_lfsr_jump(self.mask, self._degree, self._state, 1))
        return bit
