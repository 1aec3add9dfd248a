"""Bit -> symbol FSM encoder on the GPU -- mirrors waveforms/cpm/trellis/encoder.py:7-59.

``TrellisEncoder(trellis)(bits)`` accepts a uint8 CUDA tensor of 0/1 values, shape
[N] (one frame, the reference's form) or [B, N] (batch of independent frames), and
returns int8 symbols.  The encoder position ``i`` and ``state`` persist across calls
as in the reference (encoder.py:11-12,44-46); for batches every frame carries its own.
The four SOQPSK/memoryless reference trellises run through closed-form kernels
(prefix-XOR + stencil, wf_bits_encode); any other Trellis runs table-driven
(wf_trellis_encode)."""
from __future__ import annotations

import ctypes as C

import torch

from ... import _dev, _ffi
from . import model as _m

_CLOSED_FORM = (
    (_m.SOQPSKTrellis4x2DiffEncoded, _ffi.ENC_SOQPSK_DIFF),
    (_m.SOQPSKTrellis4x2, _ffi.ENC_SOQPSK),
    (_m.SimpleTrellis2, _ffi.ENC_BINARY),
    (_m.SimpleTrellis4, _ffi.ENC_QUATERNARY),
)


class TrellisEncoder:
    def __init__(self, trellis: _m.Trellis) -> None:
        self.trellis = trellis
        self.input_cardinality = trellis.input_cardinality
        self.i = 0
        self.state = 0
        self.forward_branch_mapping = [
            [_m.forward_map(st, column) for st in range(trellis.states)] for column in trellis.branches
        ]
        self._enc_id = next((eid for ref, eid in _CLOSED_FORM if _m.same_trellis(trellis, ref)), None)
        self._tables = None if self._enc_id is not None else _m.flatten_forward(trellis)
        self._batch_state = None  # int32[B, 2] on the device for batched use

    def _state_tensor(self, batch: int, device, was_1d: bool) -> torch.Tensor:
        if was_1d:
            return torch.tensor([[self.i, self.state]], dtype=torch.int32, device=device)
        if self._batch_state is None or self._batch_state.shape[0] != batch or self._batch_state.device != device:
            self._batch_state = torch.tensor([[self.i, self.state]], dtype=torch.int32, device=device).repeat(batch, 1)
        return self._batch_state

    def encode(self, bits: torch.Tensor) -> torch.Tensor:
        bits = _dev.require_cuda(bits, "bits", torch.uint8)
        frames, was_1d = _dev.as_batch(bits)
        B, N = frames.shape
        if N % self.input_cardinality:
            msg = "Input length must be a multiple of FSM cardinality."
            raise ValueError(msg)
        S = N // self.input_cardinality
        syms = torch.empty((B, S), dtype=torch.int8, device=frames.device)
        st = self._state_tensor(B, frames.device, was_1d)
        lib = _ffi.lib()
        with torch.cuda.device(frames.device):
            if self._enc_id is not None:
                rc = lib.wf_bits_encode(_dev.ptr(frames), B, N, self._enc_id, _dev.ptr(st), None, 0, _dev.ptr(syms),
                                        _dev.stream_ptr(frames.device))
            else:
                nxt, out = self._tables
                rc = lib.wf_trellis_encode(_dev.ptr(frames), B, N, nxt.ctypes.data_as(C.c_void_p),
                                           out.ctypes.data_as(C.c_void_p), self.trellis.columns, self.trellis.states,
                                           self.input_cardinality, _dev.ptr(st), _dev.ptr(syms),
                                           _dev.stream_ptr(frames.device))
        _ffi.check(rc)
        if was_1d:
            self.i, self.state = (int(v) for v in st[0].tolist())
            return syms[0]
        return syms

    def __call__(self, bits: torch.Tensor) -> torch.Tensor:
        return self.encode(bits)
