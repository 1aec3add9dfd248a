"""Complex AWGN on the GPU -- mirrors waveforms/noise.py:5-32.

The reference draws numpy PCG64 normals; here the generator is counter based
(Philox4x32-10 + Box-Muller inside wf_awgn_fill/wf_awgn_add): sample n of global frame f
at Eb/N0 point p under seed s is a pure function of (s, p, f, n), so results do not
depend on batch size, chunking or the number of GPUs.  ``rng`` is a ``PhiloxRNG``
(seed + running frame counter); the module default has seed 1 like noise.DEFAULT_RNG."""
from __future__ import annotations

import torch

from . import _dev, _ffi


class PhiloxRNG:
    """Stateful handle over the stateless generator: every call consumes `batch` frame ids."""

    def __init__(self, seed: int = 1, first_frame: int = 0) -> None:
        self.seed = int(seed)
        self.next_frame = int(first_frame)

    def take(self, batch: int) -> int:
        first = self.next_frame
        self.next_frame += int(batch)
        return first


DEFAULT_RNG = PhiloxRNG(seed=1)


def generate_complex_awgn(sigma, size: int, rng: PhiloxRNG | None = None, *, batch: int | None = None,
                          device=None) -> torch.Tensor:
    """``size`` complex samples with standard deviation ``sigma`` per real component
    (noise.py:8-32).  Returns complex64[size]; with ``batch`` given, complex64[batch, size]
    and ``sigma`` may be a float32 CUDA tensor of per-frame values."""
    rng = rng or DEFAULT_RNG
    dev = torch.device(device) if device is not None else _dev.default_device()
    B = 1 if batch is None else int(batch)
    if isinstance(sigma, torch.Tensor):
        sig = _dev.require_cuda(sigma, "sigma").to(torch.float32).reshape(-1)
        if sig.numel() == 1:
            sig = sig.repeat(B)
        dev = sig.device
    else:
        sig = torch.full((B,), float(sigma), dtype=torch.float32, device=dev)
    if sig.numel() != B:
        raise ValueError("sigma must be a scalar or have one entry per frame")
    out = torch.empty((B, int(size), 2), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _ffi.check(_ffi.lib().wf_awgn_fill(_dev.ptr(out), B, int(size), _dev.ptr(sig), None,
                                           rng.seed & (2**64 - 1), rng.take(B), None, _dev.stream_ptr(dev)))
    z = torch.view_as_complex(out)
    return z[0] if batch is None else z


def add_awgn(signal: torch.Tensor, sigma: torch.Tensor, rng: PhiloxRNG | None = None, *, point=None,
             first_frame: int | None = None, frame_ids: torch.Tensor | None = None) -> torch.Tensor:
    """In-place ``signal[b] += sigma[b] * (N(0,1) + j N(0,1))`` on complex[B, T] (wf_awgn_add).
    ``frame_ids`` (int64 CUDA tensor [B], optional) gives every row its own global frame id."""
    rng = rng or DEFAULT_RNG
    # contiguity is checked on the caller's tensor: a strided view would otherwise be copied and the
    # "in-place" add would land in the copy
    if isinstance(signal, torch.Tensor) and (signal.dim() != 2 or not signal.is_contiguous()):
        raise ValueError("signal must be a contiguous [B, T] complex tensor")
    signal = _dev.require_cuda(signal, "signal")
    B, T = signal.shape
    sig = _dev.require_cuda(sigma, "sigma", torch.float32)
    if point is not None:
        point = _dev.require_cuda(point, "point", torch.int32)
    first = rng.take(B) if first_frame is None else int(first_frame)
    if frame_ids is not None:
        frame_ids = _dev.require_cuda(frame_ids, "frame_ids", torch.int64)
    with torch.cuda.device(signal.device):
        _ffi.check(_ffi.lib().wf_awgn_add(_dev.ptr(signal), _dev.COMPLEX[signal.dtype], B, T, T, _dev.ptr(sig),
                                          _dev.ptr(point), rng.seed & (2**64 - 1), first, _dev.ptr(frame_ids),
                                          _dev.stream_ptr(signal.device)))
    return signal


def sigma_from_ebn0_db(ebn0_db, sps: int, bits_per_symbol: int = 1):
    """Invert examples/soqpsk_detection.py:132, ebn0 = 10 log10(sps / (2 sigma^2)), for a
    unit-amplitude signal carrying ``bits_per_symbol`` bits per symbol."""
    import numpy as np

    return np.sqrt(sps / (2.0 * bits_per_symbol * 10.0 ** (np.asarray(ebn0_db, dtype=np.float64) / 10.0)))
