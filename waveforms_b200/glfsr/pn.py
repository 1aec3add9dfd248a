"""PN-n maximal-length sequences -- mirrors waveforms/glfsr/pn.py:6-107.

The XAPP052 tap table lives in the C library (wf_lfsr_mask); GALOIS_LFSR_POLYS is
rebuilt from the masks so that code indexing it keeps working."""
from __future__ import annotations

import torch

from .. import _dev, _ffi
from .glfsr import GLFSR


def generate_mask(degree: int) -> int:
    """Polynomial mask of the PN-``degree`` Galois LFSR; KeyError outside 2..64 (pn.py:75-90)."""
    if not isinstance(degree, int) or not 1 < degree < 65:
        err = f"PRBS Polynomial Not Defined for {degree}."
        raise KeyError(err)
    return int(_ffi.lib().wf_lfsr_mask(degree))


def _taps(mask: int) -> list[int]:
    return [b + 1 for b in range(63, -1, -1) if (mask >> b) & 1]


class _PolyTable:
    """Lazy list-like view: GALOIS_LFSR_POLYS[degree] -> tap list, entries 0/1 empty."""

    def __len__(self) -> int:
        return 65

    def __getitem__(self, degree: int) -> list[int]:
        if degree in (0, 1):
            return []
        if not 1 < degree < 65:
            raise IndexError(degree)
        return _taps(generate_mask(degree))


GALOIS_LFSR_POLYS = _PolyTable()


class PNSequence(GLFSR):
    def __init__(self, degree: int) -> None:
        self.degree = degree
        super().__init__(generate_mask(degree), (1 << degree) - 1)

    @property
    def period(self) -> int:
        return (1 << self.degree) - 1

    def generate_sequence(self) -> list[int]:
        """The next 2^degree - 1 bits as a Python list (pn.py:98-107)."""
        return self.next_bits(self.period).cpu().tolist()

    def frames(self, first_frame: int, batch: int, nbits: int, device=None, packed: bool = False) -> torch.Tensor:
        """Frames f = first_frame .. first_frame+batch-1 of the periodic stream started from the
        all-ones state: frame f = stream bits [f*nbits, (f+1)*nbits).  Does not advance the object.
        Returns uint8[batch, nbits] (or uint32[batch, ceil(nbits/32)] when ``packed``)."""
        dev = torch.device(device) if device is not None else _dev.default_device()
        ldw = (nbits + 31) // 32
        bits = None if packed else torch.empty((batch, nbits), dtype=torch.uint8, device=dev)
        words = torch.zeros((batch, ldw), dtype=torch.int32, device=dev) if packed else None
        with torch.cuda.device(dev):
            _ffi.check(_ffi.lib().wf_prbs_encode(self.mask, self.degree, self.period, self.period, None, first_frame,
                                                 batch, nbits, _ffi.ENC_BINARY, None, _dev.ptr(bits), _dev.ptr(words),
                                                 ldw, None, _dev.stream_ptr(dev)))
        return words if packed else bits
