"""Phase tree -- waveforms/viz/tree.py:44-146 without the figure."""
from __future__ import annotations

import numpy as np
import torch

from .. import _dev, _ffi
from ..cpm.modulate import modulate_frames


def plot_phase_tree(signal: torch.Tensor, sps: int, off: float | None = None, modulo: int = 4, color=None, axis=None):
    """tree.py:64-71: per chunk of modulo * sps samples np.unwrap(np.angle(chunk)) - (off or the chunk's first
    value).  signal complex64 CUDA [T].  Returns (t_array float64 numpy [modulo*sps], phases float32 CUDA
    [n_chunks, modulo*sps])."""
    signal = _dev.require_cuda(signal, "signal", torch.complex64)
    rows, was_1d = _dev.as_batch(signal)
    B, T = rows.shape
    span = int(sps) * int(modulo)
    n_chunks = T // span
    out = torch.empty((B, n_chunks, span), dtype=torch.float32, device=rows.device)
    with torch.cuda.device(rows.device):
        _ffi.check(_ffi.lib().wf_phase_tree(_dev.ptr(rows), B, T, T, span, int(off is not None), float(off or 0.0),
                                            _dev.ptr(out), _dev.stream_ptr(rows.device)))
    t_array = np.linspace(0, modulo, span, endpoint=False)
    return t_array, (out[0] if was_1d else out)


def generate_cpm_phase_tree(pulse_filter, mod_index, encoder, sps: int, axis=None):
    """tree.py:89-146: every input sequence of bps * L bits behind an all-zero prefix of the same length is
    encoded from state 0, duplicate symbol sequences are dropped (sorted), every survivor is modulated -- ONE
    batched wf_cpm_modulate launch instead of a Python loop -- and samples [L sps - 1, 2 L sps - 1) of each form
    one branch of the tree.  Returns (t_array, phases float32 CUDA [n_sequences, L * sps])."""
    g = np.asarray(pulse_filter, dtype=np.float64)
    bps = encoder.input_cardinality
    length = int(g.size // sps)
    packed_len = bps * length
    n = 2 ** packed_len
    bits = np.zeros((n, 2 * packed_len), dtype=np.uint8)
    for j in range(packed_len):
        bits[:, packed_len + j] = (np.arange(n) >> j) & 1
    dev = _dev.default_device()
    # every row starts from (encoder.i, state 0): the reference resets .state per sequence (tree.py:118) while .i
    # keeps running -- by 2 L symbols per sequence, an even number, so the trellis column parity never changes
    encoder.state = 0
    encoder._batch_state = None
    saved_i = encoder.i
    syms = encoder.encode(torch.from_numpy(bits).to(dev))        # [n, 2 L]
    encoder._batch_state = None
    encoder.i = saved_i + n * (2 * packed_len // bps)
    encoder.state = 0
    uniq = np.unique(syms.cpu().numpy(), axis=0)                 # sorted(set(tuples)): lexicographic, as np.unique sorts
    h = [float(mod_index)] if np.isscalar(mod_index) else mod_index
    sig = modulate_frames(torch.from_numpy(uniq).to(dev), h, g, sps)
    lo = length * sps - 1
    plotted = torch.zeros(uniq.shape[0] * length * sps + 1, dtype=torch.complex64, device=dev)
    plotted[:-1] = sig[:, lo: lo + length * sps].reshape(-1)
    return plot_phase_tree(plotted, sps, modulo=length)
