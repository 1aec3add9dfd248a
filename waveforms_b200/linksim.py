"""Batched Monte-Carlo link simulation: the headless, batched equivalent of
examples/soqpsk_detection.py:35-216 (bits -> precode -> CPM modulate -> AWGN ->
matched-filter bank -> trellis detection -> error count), one kernel per stage over a
whole batch of frames, counters left on the device.

    PRBS+precode (wf_prbs_encode) -> modulate+AWGN (wf_cpm_modulate, fused Philox noise)
        -> decimated PAM/PT bank (wf_mf_bank) -> Viterbi + error count (wf_viterbi_soqpsk)

Frames are independent, so multi-GPU runs shard frames across ranks (one process per
GPU) and all-reduce only the int64 counters (``LinkSimulator.sweep``).  Frame f of
every Eb/N0 point carries PN bits [f*N, (f+1)*N) of the periodic PN stream and noise
keyed by (seed, point, f): results are identical for any batch size, chunking or
number of GPUs.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field

import numpy as np
import torch

from . import _dev, _ffi
from . import mfbank as _mf
from .cpm.multih import MULTIH_IRIG_DENOM, MULTIH_IRIG_NUMER, freq_pulse_multih_irig
from .cpm.pcmfm import PCMFM_DENOM, PCMFM_NUMER, freq_pulse_pcmfm
from .cpm.soqpsk import (SOQPSK_DENOM, SOQPSK_NUMER, freq_pulse_soqpsk_a, freq_pulse_soqpsk_b, freq_pulse_soqpsk_mil,
                         freq_pulse_soqpsk_tg)
from .glfsr.pn import generate_mask
from .noise import sigma_from_ebn0_db

# (encoder id, mod indices, pulse designer, bits per symbol) -- examples/irig_comparison.py:50-72
WAVEFORMS = {
    "soqpsk-tg": (_ffi.ENC_SOQPSK_DIFF, [SOQPSK_NUMER / SOQPSK_DENOM], freq_pulse_soqpsk_tg, 1),
    "soqpsk-mil": (_ffi.ENC_SOQPSK_DIFF, [SOQPSK_NUMER / SOQPSK_DENOM], freq_pulse_soqpsk_mil, 1),
    "soqpsk-a": (_ffi.ENC_SOQPSK_DIFF, [SOQPSK_NUMER / SOQPSK_DENOM], freq_pulse_soqpsk_a, 1),
    "soqpsk-b": (_ffi.ENC_SOQPSK_DIFF, [SOQPSK_NUMER / SOQPSK_DENOM], freq_pulse_soqpsk_b, 1),
    "pcmfm": (_ffi.ENC_BINARY, [PCMFM_NUMER / PCMFM_DENOM], freq_pulse_pcmfm, 1),
    "multih": (_ffi.ENC_QUATERNARY, list(MULTIH_IRIG_NUMER / MULTIH_IRIG_DENOM), freq_pulse_multih_irig, 2),
}


@dataclass
class LinkConfig:
    waveform: str = "soqpsk-tg"
    n_bits: int = 4096          # bits per frame
    sps: int = 8
    detector: str = "PAM"       # "PAM" or "PT" (examples/soqpsk_detection.py:177)
    timing_offset: int = 0      # soqpsk_detection.py:184-187 sign convention: (n + offset) % sps == 0
    length: int = 2             # detector window
    differential: bool = True
    pn_degree: int = 15
    seed: int = 1
    dtype: torch.dtype = torch.complex64
    extra: dict = field(default_factory=dict)


class LinkSimulator:
    """Owns the work buffers for up to ``max_batch`` frames and runs the four-kernel chain."""

    def __init__(self, cfg: LinkConfig | None = None, device=None, max_batch: int = 4096) -> None:
        self.cfg = cfg = cfg or LinkConfig()
        self.device = torch.device(device) if device is not None else _dev.default_device()
        self._max_batch_arg = int(max_batch)
        if cfg.waveform not in WAVEFORMS:
            raise KeyError(f"unknown waveform {cfg.waveform!r}; choose from {sorted(WAVEFORMS)}")
        enc, h, pulse_fn, bps = WAVEFORMS[cfg.waveform]
        self.soqpsk = cfg.waveform.startswith("soqpsk")
        self.encoder = (enc if cfg.differential else _ffi.ENC_SOQPSK) if self.soqpsk else enc
        self.h = _dev.host_f64(h)
        self.bits_per_symbol = bps
        self.pulse = _dev.host_f64(pulse_fn(cfg.sps))
        self.mask = generate_mask(cfg.pn_degree)
        self.period = (1 << cfg.pn_degree) - 1
        self.n_sym = cfg.n_bits // bps
        self.n_samples = (self.n_sym + 1) * cfg.sps
        if not self.soqpsk:
            self._init_generic(h, bps)
            return
        self.first_instant, self.n_calls = _mf.symbol_instants(self.n_samples, cfg.sps, cfg.timing_offset, cfg.length)
        self.n_compared = max(0, min(self.n_sym, self.n_calls - cfg.length))
        if cfg.detector.upper() == "PAM":
            taps, comb = _mf.pam_taps(self.pulse, float(self.h[0]), cfg.sps)
        elif cfg.detector.upper() == "PT":
            taps, comb = _mf.pt_taps(self.pulse, float(self.h[0]), cfg.sps), None
        else:
            raise ValueError("detector must be 'PAM' or 'PT'")
        self._taps = np.asarray(taps)
        self._tap_re = _dev.host_f64(self._taps.real.ravel())
        self._tap_im = _dev.host_f64(self._taps.imag.ravel()) if np.iscomplexobj(self._taps) else None
        self._comb_re = None if comb is None else _dev.host_f64(np.asarray(comb).real.ravel())
        self._comb_im = None if comb is None else _dev.host_f64(np.asarray(comb).imag.ravel())
        self._n_out = 3
        # PAM bank + length-2 detector: the bank stores its two raw filter outputs (16 B per call)
        # and the detector applies the 3x2 pseudo-symbol combine itself (wf_viterbi_soqpsk_combined):
        # same counters bit for bit, a third less traffic between the two kernels.
        # extra["combine_in_bank"] = True keeps the reference-shaped [B, calls, 3] array instead.
        self.raw_bank = (comb is not None and cfg.length == 2 and self._taps.shape[0] == 2
                         and not bool(cfg.extra.get("combine_in_bank", False)))
        if self.raw_bank:
            self._n_out = 2
        # Fused modulate -> AWGN -> PAM bank (wf_cpm_modulate_bank): the noisy baseband never reaches HBM and
        # the rx buffer is not allocated.  Same z, bit for bit, as the two-kernel path, and faster (6.9 against
        # 4.2 + 3.8 ms per 65536 frames), so it is the DEFAULT whenever the library covers the shape (SOQPSK-TG /
        # -A at sps 8, PAM bank, offset 0); extra["fused"] = False keeps wf_cpm_modulate + wf_mf_bank.
        self.fused = bool(cfg.extra.get("fused", True)) and self.raw_bank and cfg.dtype == torch.complex64 and bool(
            _ffi.lib().wf_cpm_modulate_bank_supported(int(self.pulse.size), cfg.sps, int(self._taps.shape[0]),
                                                      int(self._taps.shape[1]), (int(self._taps.shape[1]) - 1) // 2,
                                                      int(self.first_instant), int(self.n_sym), int(self.n_calls)))
        # which fused kernel: 0 = one block per frame, 1 / 2 = one warp per frame with 64- / 128-interval chunks
        self.fused_tune = int(cfg.extra.get("fused_tune", 1))
        self.one_call = bool(cfg.extra.get("one_call", True))  # False: three C calls instead of wf_link_sim_fused
        self.stage_names = (("prbs_encode", "modulate_awgn_bank", "viterbi_bert") if self.fused
                            else ("prbs_encode", "modulate_awgn", "mf_bank", "viterbi_bert"))
        self.max_batch = 0
        self._alloc(int(max_batch))
        self.launches = 0  # kernels launched so far (len(stage_names) per run_batch)

    def _init_generic(self, h, bps) -> None:
        """PCM/FM and multi-h: generic pulse-truncated trellis detector (no reference detector exists)."""
        from fractions import Fraction

        from .viterbi.cpm import CPMTrellisDetector

        cfg = self.cfg
        fr = [Fraction(float(x)).limit_denominator(64) for x in h]
        den = int(np.lcm.reduce([f.denominator for f in fr]))
        self.det = CPMTrellisDetector(self.pulse, [int(f * den) for f in fr], den, 2 ** bps, cfg.sps,
                                      memory=int(cfg.extra.get("memory", 2)), depth=int(cfg.extra.get("depth", 0)))
        self.skip = int(cfg.extra.get("skip", 4))
        self.first_instant, self.n_calls = self.det.first_instant, self.n_sym
        self.n_compared = max(0, self.n_sym - 2 * self.skip) * bps  # compared BITS per frame, as for SOQPSK
        self._taps = self.det.taps
        self._tap_re = _dev.host_f64(self._taps.real.ravel())
        self._tap_im = _dev.host_f64(self._taps.imag.ravel())
        self._comb_re = self._comb_im = None
        self._n_out = self._taps.shape[0]
        self.fused = False
        # wf_cpm_detect: the hypothesis bank runs inside the detector kernel (no bank launch, no z array);
        # taken whenever the library covers the shape (the two IRIG waveforms at sps 8, complex64);
        # extra["unfused_detector"] = True keeps wf_mf_bank + wf_viterbi_cpm (the generic, parity path)
        T = self.n_samples
        self.detect_fused = (not bool(cfg.extra.get("unfused_detector", False))) and self.det.supports_fused(
            cfg.dtype, T, (T + 1) & ~1, self.n_calls)
        self.tune = int(cfg.extra.get("tune", 4))  # path metrics re-normalised every 4 steps (1 = every step, as the oracle)
        self.stage_names = (("prbs_encode", "modulate_awgn", "detect_fused") if self.detect_fused
                            else ("prbs_encode", "modulate_awgn", "mf_bank", "viterbi_bert"))
        self.max_batch = 0
        self._alloc(int(self._max_batch_arg))
        self.launches = 0

    # ------------------------------------------------------------------ buffers
    def _alloc(self, max_batch: int) -> None:
        cfg, dev = self.cfg, self.device
        self.max_batch = max_batch
        self.ld_words = (cfg.n_bits + 31) // 32
        self.ld_rx = (self.n_samples + 1) & ~1
        self.ld_z = (self.n_calls + 3) & ~3 if self.soqpsk else self.n_calls
        self._ws = None
        if self.soqpsk and self.fused:
            # ONE workspace in the layout of wf_link_sim_fused (packed bits | symbols | raw bank outputs, each
            # 256-byte aligned); packed / syms / z are views of it, so the one-call path (run_batch without
            # per-kernel events) and the three-call path share the memory
            self._link_cfg = self._make_link_cfg()
            total = int(_ffi.lib().wf_link_sim_workspace_bytes(C.byref(self._link_cfg), max_batch))
            if total < 0:
                _ffi.check(total)
            up = lambda v: (v + 255) // 256 * 256  # noqa: E731
            n_p, n_s = max_batch * self.ld_words * 4, max_batch * self.n_sym
            n_z = max_batch * self.ld_z * self._n_out * 8
            o_s = up(n_p)
            o_z = o_s + up(n_s)
            assert o_z + up(n_z) == total, (o_z + up(n_z), total)
            self._ws = torch.empty(max(total, 1), dtype=torch.uint8, device=dev)
            self.packed = self._ws[:n_p].view(torch.int32).view(max_batch, self.ld_words)
            self.syms = self._ws[o_s:o_s + n_s].view(torch.int8).view(max_batch, self.n_sym)
            self.z = self._ws[o_z:o_z + n_z].view(cfg.dtype).view(max_batch, self.ld_z, self._n_out)
            self.rx = None
            return
        self.packed = torch.zeros((max_batch, self.ld_words), dtype=torch.int32, device=dev)
        self.syms = torch.empty((max_batch, self.n_sym), dtype=torch.int8, device=dev)
        self.rx = None if self.fused else torch.empty((max_batch, self.ld_rx), dtype=cfg.dtype, device=dev)
        no_z = (not self.soqpsk) and self.detect_fused
        self.z = None if no_z else torch.empty((max_batch, self.ld_z, self._n_out), dtype=cfg.dtype, device=dev)

    def _make_link_cfg(self):
        """struct wf_link_cfg for this simulator (the host tables it points to are kept alive by self)."""
        cfg = self.cfg
        nt = int(self._taps.shape[1])
        return _ffi.LinkCfg(
            mask=self.mask, init_state=self.period, period=self.period, degree=cfg.pn_degree, encoder=self.encoder,
            n_bits=cfg.n_bits, mod_index=_dev.f64_ptr(self.h), pulse=_dev.f64_ptr(self.pulse), n_h=int(self.h.size),
            K=int(self.pulse.size), sps=cfg.sps, tune=self.fused_tune, initial_phase=math.pi / 4, derotate=-math.pi / 4,
            seed=cfg.seed & (2**64 - 1), taps_re=_dev.f64_ptr(self._tap_re), n_filt=int(self._taps.shape[0]), n_taps=nt,
            center=(nt - 1) // 2, diff=int(cfg.differential), first_instant=int(self.first_instant),
            n_calls=int(self.n_calls), combine_re=_dev.f64_ptr(self._comb_re), combine_im=_dev.f64_ptr(self._comb_im))

    def workspace_bytes(self) -> int:
        if self._ws is not None:
            return self._ws.numel()
        return sum(t.numel() * t.element_size() for t in (self.packed, self.syms, self.rx, self.z) if t is not None)

    # ------------------------------------------------------------------ one batch
    def run_batch(self, first_frame: int, sigma: torch.Tensor, point: torch.Tensor | None, counters: torch.Tensor,
                  events: list | None = None, frame_ids: torch.Tensor | None = None,
                  bit_offsets: torch.Tensor | None = None) -> None:
        """Frames first_frame .. first_frame+B-1 (B = len(sigma) <= max_batch) through the whole
        chain; adds (bit errors, symbol errors, compared bits) to counters[point[b]].  Everything is
        enqueued on the current stream; nothing is synchronised.  ``events`` (optional list)
        receives len(stage_names) + 1 CUDA events recorded on that stream around the kernels.
        ``frame_ids`` / ``bit_offsets`` (int64[B] device tensors, both or neither): row b is global
        frame frame_ids[b], whose PN bits start bit_offsets[b] = frame_ids[b] * n_bits (mod period)
        into the stream -- how ``run`` packs frames of several Eb/N0 points into one launch."""
        cfg, lib, dev = self.cfg, _ffi.lib(), self.device
        B = sigma.shape[0]
        if B > self.max_batch:
            raise ValueError(f"batch {B} exceeds max_batch {self.max_batch}")
        st = _dev.stream_ptr(dev)
        dt = _dev.COMPLEX[cfg.dtype]
        seed = cfg.seed & (2**64 - 1)
        raw = self.soqpsk and self.raw_bank
        if (frame_ids is None) != (bit_offsets is None):
            raise ValueError("frame_ids and bit_offsets go together")
        fid, boff = _dev.ptr(frame_ids), _dev.ptr(bit_offsets)

        def mark():
            if events is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record(torch.cuda.current_stream(dev))
                events.append(e)

        with torch.cuda.device(dev):
            if self._ws is not None and events is None and self.one_call:
                # the whole chain in ONE C call (wf_link_sim_fused): same three kernels, same counters
                _ffi.check(lib.wf_link_sim_fused(C.byref(self._link_cfg), B, int(first_frame), fid, boff, _dev.ptr(sigma),
                                                 _dev.ptr(point), _dev.ptr(self._ws), self._ws.numel(), _dev.ptr(counters), st))
                self.launches += 3
                return
            mark()
            _ffi.check(lib.wf_prbs_encode(self.mask, cfg.pn_degree, self.period, self.period, boff, int(first_frame), B,
                                          cfg.n_bits, self.encoder, None, None, _dev.ptr(self.packed), self.ld_words,
                                          _dev.ptr(self.syms), st))
            mark()
            if self.fused:
                _ffi.check(lib.wf_cpm_modulate_bank(_dev.ptr(self.syms), B, self.n_sym, _dev.f64_ptr(self.h), self.h.size,
                                                    _dev.f64_ptr(self.pulse), self.pulse.size, cfg.sps, math.pi / 4,
                                                    -math.pi / 4, _dev.ptr(sigma), _dev.ptr(point), seed, int(first_frame),
                                                    fid, _dev.f64_ptr(self._tap_re), self._taps.shape[0], self._taps.shape[1],
                                                    (self._taps.shape[1] - 1) // 2, self.first_instant, self.n_calls,
                                                    _dev.ptr(self.z), self.ld_z, self.fused_tune, st))
                mark()
                _ffi.check(lib.wf_viterbi_soqpsk_combined(_dev.ptr(self.z), dt, B, self.n_calls, self.ld_z, 2,
                                                          _dev.f64_ptr(self._comb_re), _dev.f64_ptr(self._comb_im),
                                                          int(cfg.differential), _dev.ptr(self.packed), self.ld_words,
                                                          self.n_sym, _dev.ptr(point), _dev.ptr(counters), st))
                mark()
                self.launches += 3
                return
            _ffi.check(lib.wf_cpm_modulate(_dev.ptr(self.syms), B, self.n_sym, _dev.f64_ptr(self.h), self.h.size,
                                           _dev.f64_ptr(self.pulse), self.pulse.size, cfg.sps, math.pi / 4, -math.pi / 4,
                                           dt, _dev.ptr(self.rx), self.ld_rx, _dev.ptr(sigma), _dev.ptr(point), seed,
                                           int(first_frame), fid, st))
            mark()
            if not self.soqpsk and self.detect_fused:
                d = self.det
                _ffi.check(lib.wf_cpm_detect(_dev.ptr(self.rx), dt, B, self.n_samples, self.ld_rx, _dev.f64_ptr(self._tap_re),
                                             _dev.f64_ptr(self._tap_im), cfg.sps, self.first_instant, self.n_calls, d.M, d.Lp,
                                             d.p, d.h_num.ctypes.data, d.n_h, d.depth, None, _dev.ptr(self.syms), self.n_sym,
                                             self.skip, self.skip, _dev.ptr(point), _dev.ptr(counters), self.tune, st))
                mark()
                self.launches += 3
                return
            _ffi.check(lib.wf_mf_bank(_dev.ptr(self.rx), dt, B, self.n_samples, self.ld_rx, _dev.f64_ptr(self._tap_re),
                                      None if self._tap_im is None else _dev.f64_ptr(self._tap_im),
                                      self._taps.shape[0], self._taps.shape[1],
                                      (self._taps.shape[1] - 1) // 2 if self.soqpsk else cfg.sps - 1,
                                      None if (self._comb_re is None or raw) else _dev.f64_ptr(self._comb_re),
                                      None if (self._comb_im is None or raw) else _dev.f64_ptr(self._comb_im), self._n_out,
                                      cfg.sps, self.first_instant, self.n_calls, _dev.ptr(self.z), self.ld_z, st))
            mark()
            if not self.soqpsk:
                d = self.det
                _ffi.check(lib.wf_viterbi_cpm(_dev.ptr(self.z), dt, B, self.n_calls, self.ld_z, d.M, d.Lp, d.p,
                                              d.h_num.ctypes.data, d.n_h, d.depth, None, _dev.ptr(self.syms), self.n_sym,
                                              self.skip, self.skip, _dev.ptr(point), _dev.ptr(counters), st))
                mark()
                self.launches += 4
                return
            if raw:
                _ffi.check(lib.wf_viterbi_soqpsk_combined(_dev.ptr(self.z), dt, B, self.n_calls, self.ld_z, 2,
                                                          _dev.f64_ptr(self._comb_re), _dev.f64_ptr(self._comb_im),
                                                          int(cfg.differential), _dev.ptr(self.packed), self.ld_words,
                                                          self.n_sym, _dev.ptr(point), _dev.ptr(counters), st))
                mark()
                self.launches += 4
                return
            _ffi.check(lib.wf_viterbi_soqpsk(_dev.ptr(self.z), dt, B, self.n_calls, self.ld_z, int(cfg.differential),
                                             cfg.length, None, None, _dev.ptr(self.packed), self.ld_words, None,
                                             self.n_sym, _dev.ptr(point), _dev.ptr(counters), None, st))
            mark()
        self.launches += 4

    # ------------------------------------------------------------------ host-facing calls
    def _staging(self, slot: int):
        """Pinned host tables + their device twins for one batch, allocated once (two slots: the
        host fills slot k+1 while the copy of slot k may still be in flight; an event per slot
        says when its pinned memory may be overwritten)."""
        if not hasattr(self, "_stage"):
            self._stage = []
            for _ in range(2):
                h = {"sigma": torch.empty(self.max_batch, dtype=torch.float32).pin_memory(),
                     "point": torch.empty(self.max_batch, dtype=torch.int32).pin_memory(),
                     "frame": torch.empty(self.max_batch, dtype=torch.int64).pin_memory(),
                     "off": torch.empty(self.max_batch, dtype=torch.int64).pin_memory()}
                d = {k: torch.empty_like(v, device=self.device) for k, v in h.items()}
                self._stage.append((h, d, torch.cuda.Event()))
            self._stage_used = [False, False]
        return self._stage[slot]

    def _bit_offsets(self, frames: np.ndarray) -> np.ndarray:
        """(frame * n_bits) mod PN period, exact for any frame id."""
        n = self.cfg.n_bits
        if self.period < (1 << 31):
            return ((frames % self.period) * (n % self.period)) % self.period
        return np.array([(int(f) * n) % self.period for f in frames], dtype=np.int64)

    def run(self, ebn0_db, n_frames: int, first_frame: int = 0, active=None) -> np.ndarray:
        """End-to-end call with HOST inputs and outputs: ``ebn0_db`` is a host sequence of P
        Eb/N0 points; every point is simulated on frames first_frame .. first_frame+n_frames-1.
        The (point, frame) work list is cut into batches of ``max_batch`` rows REGARDLESS of point
        boundaries -- per-row sigma / point / frame-id / PN-offset tables go through preallocated
        pinned memory -- so a 13-point sweep of 96 frames is one launch of the chain, not 13.
        Returns int64[P, 3] = (bit errors, symbol errors, compared bits) as a numpy array.
        ``active`` (bool[P], optional): points that get no frames keep their row and noise stream id."""
        eb = np.atleast_1d(np.asarray(ebn0_db, dtype=np.float64))
        P = eb.size
        sig = sigma_from_ebn0_db(eb, self.cfg.sps, self.bits_per_symbol).astype(np.float32)
        counters = torch.zeros((P, 3), dtype=torch.int64, device=self.device)
        pts = np.arange(P) if active is None else np.flatnonzero(np.asarray(active, dtype=bool))
        total = pts.size * int(n_frames)
        stream = torch.cuda.current_stream(self.device)
        k = 0
        for w0 in range(0, total, self.max_batch):
            B = min(self.max_batch, total - w0)
            h, d, ev = self._staging(k & 1)
            if self._stage_used[k & 1]:
                ev.synchronize()  # the previous copy out of this pinned slot has completed
            # one point, consecutive frames (the common large-batch case): constant tables, no per-row frame ids
            one_run = w0 // n_frames == (w0 + B - 1) // n_frames
            names = ("sigma", "point")
            if one_run:
                p0 = int(pts[w0 // n_frames])
                f_first = int(first_frame) + w0 % n_frames
                h["sigma"][:B].fill_(float(sig[p0]))
                h["point"][:B].fill_(p0)
            else:
                pw, fw = work_slice(pts, int(n_frames), int(first_frame), w0, B)
                h["sigma"][:B] = torch.from_numpy(sig[pw])
                h["point"][:B] = torch.from_numpy(pw.astype(np.int32))
                h["frame"][:B] = torch.from_numpy(fw)
                h["off"][:B] = torch.from_numpy(self._bit_offsets(fw).astype(np.int64))
                names += ("frame", "off")
            for nm in names:
                d[nm][:B].copy_(h[nm][:B], non_blocking=True)
            ev.record(stream)
            self._stage_used[k & 1] = True
            if one_run:
                self.run_batch(f_first, d["sigma"][:B], d["point"][:B], counters)
            else:
                self.run_batch(0, d["sigma"][:B], d["point"][:B], counters, frame_ids=d["frame"][:B],
                               bit_offsets=d["off"][:B])
            k += 1
        return counters.cpu().numpy()

    def sweep(self, ebn0_db, n_frames: int, first_frame: int = 0, group=None, active=None) -> np.ndarray:
        """Multi-GPU ``run``: this rank simulates its contiguous share of the n_frames of every
        point, then the counters are all-reduced (NCCL, SUM, int64[P,3] -- the only traffic that
        crosses NVLink).  Without an initialised process group it equals ``run``."""
        import torch.distributed as dist

        eb = np.atleast_1d(np.asarray(ebn0_db, dtype=np.float64))
        if not (dist.is_available() and dist.is_initialized()):
            return self.run(eb, n_frames, first_frame, active)
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        lo, hi = shard_range(n_frames, rank, world)
        if hi > lo:
            local = self.run(eb, hi - lo, first_frame + lo, active)
        else:
            local = np.zeros((eb.size, 3), dtype=np.int64)
        return allreduce_counters(local, self.device, group)

    def waterfall(self, ebn0_db, min_errors: int = 100, max_bits: float = 1e11, chunk_frames: int | None = None,
                  first_frame: int = 0, group=None) -> np.ndarray:
        """BER waterfall with a stop rule (BASELINE configs[3]): every Eb/N0 point is
        simulated in chunks of ``chunk_frames`` frames (default: one full batch per rank) until it
        has collected ``min_errors`` bit errors or ``max_bits`` compared bits.  Chunks are sharded
        over the ranks of ``group`` exactly like ``sweep``; the counter all-reduce after each chunk
        gives every rank the same global view, so all ranks retire a point at the same time.  Frame
        ids advance by ``chunk_frames`` per chunk for every point, and a point's noise stream is
        keyed by its index in ``ebn0_db``, so the result does not depend on the number of GPUs."""
        import torch.distributed as dist

        eb = np.atleast_1d(np.asarray(ebn0_db, dtype=np.float64))
        world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        chunk = int(chunk_frames) if chunk_frames else world * self.max_batch
        if chunk <= 0 or self.n_compared <= 0:
            raise ValueError("chunk_frames and the compared bits per frame must be positive")
        total = np.zeros((eb.size, 3), dtype=np.int64)
        active = np.ones(eb.size, dtype=bool)
        frame = int(first_frame)
        while active.any():
            total += self.sweep(eb, chunk, frame, group, active)
            frame += chunk
            active = (total[:, 0] < min_errors) & (total[:, 2] < max_bits)
        return total


def work_slice(points: np.ndarray, n_frames: int, first_frame: int, w0: int, count: int):
    """Rows w0 .. w0+count-1 of the (point, frame) work list of ``LinkSimulator.run``: point-major,
    frames first_frame .. first_frame+n_frames-1 inside each point.  Returns (point index per row,
    global frame id per row) as int64 arrays."""
    w = np.arange(w0, w0 + count, dtype=np.int64)
    return np.asarray(points, dtype=np.int64)[w // n_frames], first_frame + (w % n_frames)


def allreduce_counters(local: np.ndarray, device, group=None) -> np.ndarray:
    """SUM all-reduce of the int64 counter table across ranks: on the GPU over NCCL
    (NVLink/NVSwitch), on host tensors when the group's backend is gloo (CPU tests)."""
    import torch.distributed as dist

    t = torch.from_numpy(np.ascontiguousarray(local, dtype=np.int64))
    if dist.get_backend(group) == "nccl":
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy()


def shard_range(n_items: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced [lo, hi) share of n_items for ``rank`` of ``world``."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def ber_confidence_interval(errors: int, bits: int, level: float = 0.95) -> tuple[float, float]:
    """Clopper-Pearson interval for a bit-error ratio (the reference computes none)."""
    from scipy.stats import beta

    if bits == 0:
        return 0.0, 1.0
    a = (1 - level) / 2
    lo = 0.0 if errors == 0 else float(beta.ppf(a, errors, bits - errors + 1))
    hi = 1.0 if errors == bits else float(beta.ppf(1 - a, errors + 1, bits - errors))
    return lo, hi
