"""CPM modulation front-end: pulses and table design on the host (numpy, set-up
time), per-frame work on the GPU.  Mirrors waveforms/cpm/ of the reference."""
