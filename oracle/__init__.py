"""CPU oracle for the waveforms hot path -- TEST INFRASTRUCTURE ONLY.

A float64 numpy/pure-Python restatement of the reference algorithm
(mcdiarmid/waveforms) for PRBS -> precode -> CPM modulate -> AWGN ->
matched-filter bank -> SOQPSK trellis detection -> error counting.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this package, and only as
the checker / the timed CPU baseline.  The product (``waveforms_b200``) never
imports it and has no CPU fallback.

Parity status: PINNED for PRBS, encoders/mappers, pulses, modulators, noise
layout, PT/PAM banks, the SOQPSK detector and the error count -- the
restatement is checked against golden vectors produced by running the
reference itself (``oracle/make_golden.py`` -> ``tests/golden/*.npz``).
UNPINNED for PCM/FM and multi-h *detection*: the reference has no detector
for them; ``oracle/cpm_generic.py`` is the builder's own float64 design.
"""
