"""Signal diagnostics -- mirror of waveforms/viz (psd.py, eye.py, constellation.py, tree.py).

The reference's functions draw matplotlib figures; figure output is unavailable on this box (no
graphics stack), so every function here returns the ARRAYS the figure is drawn from, computed by CUDA
reductions / gathers over torch CUDA tensors (csrc/diag.cu).  Names and arguments follow the
reference; ``axis`` / ``axes`` / ``color`` are accepted and ignored."""
from .constellation import constellation
from .eye import plot_eye_diagram
from .psd import plot_power_spectral_density
from .tree import generate_cpm_phase_tree, plot_phase_tree

__all__ = ["constellation", "plot_eye_diagram", "plot_power_spectral_density", "plot_phase_tree", "generate_cpm_phase_tree"]
