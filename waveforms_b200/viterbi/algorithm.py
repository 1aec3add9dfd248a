"""SOQPSK trellis detector on the GPU -- mirrors waveforms/viterbi/algorithm.py:18-101.

``SOQPSKTrellisDetector(length=2, *, differantial_encoding=True)`` keeps the reference's
constructor (including the misspelt keyword) and its streaming ``iteration(mf_outputs)``
call; ``detect(z)`` is the batched form the link simulator uses: z complex[B, calls, 3]
-> element [0] of every call's output for every frame in one kernel launch
(wf_viterbi_soqpsk), optionally with the error count of
examples/soqpsk_detection.py:200-209 fused in."""
from __future__ import annotations

import torch

from .. import _dev, _ffi
from ..cpm.trellis.model import FiniteStateMachine, SOQPSKTrellis4x2, SOQPSKTrellis4x2DiffEncoded


class SOQPSKTrellisDetector:
    def __init__(self, length: int = 2, *, differantial_encoding: bool = True) -> None:
        self.i = 0
        self.length = int(length)
        self.differential = bool(differantial_encoding)
        self.fsm = FiniteStateMachine(trellis=SOQPSKTrellis4x2DiffEncoded if differantial_encoding else SOQPSKTrellis4x2)
        self.state_exp_term = [+1j, -1, +1, -1j]
        self._state = None  # device tensor [B, 5 + 8*length]: parity, metrics[:,0], bi_history

    # ---- batched one-shot form -------------------------------------------------
    def detect(self, z: torch.Tensor, *, tx_packed: torch.Tensor | None = None, tx_symbols: torch.Tensor | None = None,
               n_symbols: int | None = None, point: torch.Tensor | None = None, counters: torch.Tensor | None = None,
               return_decisions: bool = True, n_calls: int | None = None):
        """Run a FRESH detector over every frame of z complex[B, ld>=calls, 3].

        Returns (bits uint8[B, calls], symbols int8[B, calls]) (None, None when
        ``return_decisions`` is False).  With ``counters`` (int64[P, 3] = bit errors, symbol
        errors, compared) the decisions [length:] are compared with ``tx_packed`` (uint32 words,
        LSB first) and/or ``tx_symbols`` and accumulated at row ``point[b]``."""
        z = _dev.require_cuda(z, "z")
        if z.dtype not in _dev.COMPLEX or z.dim() != 3 or z.shape[2] != 3:
            raise TypeError("z must be complex[B, calls, 3]")
        B, ld, _ = z.shape
        calls = ld if n_calls is None else int(n_calls)
        dev = z.device
        bits = syms = None
        if return_decisions:
            bits = torch.empty((B, calls), dtype=torch.uint8, device=dev)
            syms = torch.empty((B, calls), dtype=torch.int8, device=dev)
        ldw, S = 0, 0
        if tx_packed is not None:
            tx_packed = _dev.require_cuda(tx_packed, "tx_packed", torch.int32)
            ldw = tx_packed.shape[1]
            S = int(n_symbols) if n_symbols is not None else 32 * ldw
        if tx_symbols is not None:
            tx_symbols = _dev.require_cuda(tx_symbols, "tx_symbols", torch.int8)
            S = tx_symbols.shape[1] if n_symbols is None else int(n_symbols)
        if point is not None:
            point = _dev.require_cuda(point, "point", torch.int32)
        if counters is not None:
            counters = _dev.require_cuda(counters, "counters", torch.int64)
        with torch.cuda.device(dev):
            _ffi.check(_ffi.lib().wf_viterbi_soqpsk(_dev.ptr(z), _dev.COMPLEX[z.dtype], B, calls, ld,
                                                    int(self.differential), self.length, _dev.ptr(bits), _dev.ptr(syms),
                                                    _dev.ptr(tx_packed), ldw, _dev.ptr(tx_symbols), S, _dev.ptr(point),
                                                    _dev.ptr(counters), None, _dev.stream_ptr(dev)))
        return bits, syms

    # ---- streaming form: the reference's per-call interface ---------------------------
    def iterate(self, z: torch.Tensor):
        """Feed complex[calls, 3] (or [B, calls, 3]) to a detector whose state persists across
        calls; returns element [0] of each call's (bits, symbols)."""
        z = _dev.require_cuda(z, "z")
        single = z.dim() == 2
        zz = z.unsqueeze(0) if single else z
        B, calls, _ = zz.shape
        real = torch.float32 if zz.dtype == torch.complex64 else torch.float64
        if self._state is None or self._state.shape[0] != B or self._state.dtype != real:
            self._state = torch.zeros((B, 5 + 8 * self.length), dtype=real, device=zz.device)
            self._state[:, 0] = self.i & 1
        bits = torch.empty((B, calls), dtype=torch.uint8, device=zz.device)
        syms = torch.empty((B, calls), dtype=torch.int8, device=zz.device)
        with torch.cuda.device(zz.device):
            _ffi.check(_ffi.lib().wf_viterbi_soqpsk(_dev.ptr(zz), _dev.COMPLEX[zz.dtype], B, calls, calls,
                                                    int(self.differential), self.length, _dev.ptr(bits), _dev.ptr(syms),
                                                    None, 0, None, 0, None, None, _dev.ptr(self._state),
                                                    _dev.stream_ptr(zz.device)))
        self.i += calls
        return (bits[0], syms[0]) if single else (bits, syms)

    def iteration(self, mf_outputs: torch.Tensor):
        """One detector call on one matched-filter 3-vector (algorithm.py:44-101).

        The reference returns the whole ``length``-deep traceback (output, recovered_symbols)
        as float64 arrays of which its only caller keeps element [0]
        (examples/soqpsk_detection.py:196-198).  Here element [0] is the kernel's decision and
        the remaining entries -- the provisional look-ahead decisions -- are not produced: they
        are returned as NaN so that code reading beyond [0] fails visibly."""
        z = _dev.require_cuda(mf_outputs, "mf_outputs")
        if z.dtype not in _dev.COMPLEX:
            z = z.to(torch.complex128)
        b, s = self.iterate(z.reshape(1, 3))
        out = torch.full((self.length,), float("nan"), dtype=torch.float64, device=z.device)
        sym = out.clone()
        out[0] = b[0]
        sym[0] = s[0]
        return out, sym
