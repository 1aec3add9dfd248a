"""Bit / symbol error counting.  ``waveforms.bert`` is an empty package upstream; the
counting itself is inline at examples/soqpsk_detection.py:200-216.  ``count_errors``
restates those lines for CUDA tensors (wf_bert kernel); the fused form lives in the
detector (SOQPSKTrellisDetector.detect(counters=...))."""
from __future__ import annotations

import torch

from .. import _dev, _ffi


def count_errors(decisions: torch.Tensor, reference: torch.Tensor, skip: int = 2, *, dec_symbols=None,
                 ref_symbols=None, point=None, counters=None):
    """Drop the first ``skip`` detector outputs, compare the next min(len(reference),
    len(decisions) - skip) with the transmitted values, count mismatches.

    decisions / reference: uint8 [n] or [B, n] (bits).  Returns (bit_errors, symbol_errors,
    compared) as Python ints summed over the batch, or accumulates into ``counters``
    int64[P, 3] at rows ``point`` and returns it."""
    dec = _dev.require_cuda(decisions, "decisions", torch.uint8)
    ref = _dev.require_cuda(reference, "reference", torch.uint8)
    dec, _ = _dev.as_batch(dec)
    ref, _ = _dev.as_batch(ref)
    B = dec.shape[0]
    n_cmp = max(0, min(ref.shape[1], dec.shape[1] - skip))
    ds = rs = None
    if dec_symbols is not None and ref_symbols is not None:
        ds, _ = _dev.as_batch(_dev.require_cuda(dec_symbols, "dec_symbols", torch.int8))
        rs, _ = _dev.as_batch(_dev.require_cuda(ref_symbols, "ref_symbols", torch.int8))
    own = counters is None
    if own:
        counters = torch.zeros((1, 3), dtype=torch.int64, device=dec.device)
    if point is not None:
        point = _dev.require_cuda(point, "point", torch.int32)
    with torch.cuda.device(dec.device):
        _ffi.check(_ffi.lib().wf_bert(_dev.ptr(dec), _dev.ptr(ds), dec.shape[1], _dev.ptr(ref), _dev.ptr(rs),
                                      ref.shape[1], B, int(skip), n_cmp, _dev.ptr(point), _dev.ptr(counters),
                                      _dev.stream_ptr(dec.device)))
    if own:
        be, se, n = counters[0].tolist()
        return be, se, n
    return counters
