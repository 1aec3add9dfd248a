"""Receive-side filtering and timing -- mirror of waveforms/filters (lpf.py) plus the batched application
and a sampling-phase estimator that replaces the hard-coded timing offsets of
examples/soqpsk_detection.py:183-192."""
from .lpf import estimate_timing_offset, fir_filter, kaiser_fir_lpf

__all__ = ["kaiser_fir_lpf", "fir_filter", "estimate_timing_offset"]
