"""Device plumbing shared by the host-side mirror of the reference API: argument
checks (CUDA tensors only -- there is no CPU path), current-stream handle, dtype
codes.  torch is used for memory, streams and torch.distributed; all arithmetic on
frames happens in libwfb200's kernels."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _ffi

COMPLEX = {torch.complex64: _ffi.WF_F32, torch.complex128: _ffi.WF_F64}


def require_cuda(t, name: str, dtype=None):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch CUDA tensor (got {type(t).__name__}); waveforms_b200 has no CPU path")
    if not t.is_cuda:
        raise TypeError(f"{name} must live on a CUDA device; waveforms_b200 has no CPU path")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    return t.contiguous()


def stream_ptr(device=None) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t) -> C.c_void_p:
    return C.c_void_p(0) if t is None else C.c_void_p(t.data_ptr())


def as_batch(t: torch.Tensor):
    """[..] -> ([B, n], was_1d)."""
    if t.dim() == 1:
        return t.unsqueeze(0), True
    if t.dim() != 2:
        raise ValueError("expected a 1-D frame or a [batch, n] tensor")
    return t, False


def host_f64(x) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(np.atleast_1d(np.asarray(x, dtype=np.float64)))


def f64_ptr(a: np.ndarray) -> C.c_void_p:
    return a.ctypes.data_as(C.c_void_p)


def default_device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("waveforms_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())
