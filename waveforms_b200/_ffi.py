"""ctypes binding of the C ABI in include/wfb200.h (one prototype per entry point).

``lib()`` loads waveforms_b200/_lib/libwfb200.so, building it with nvcc first when it
is missing.  There is no CPU implementation behind these calls: when the library
cannot be built or loaded the import fails loudly, and on a machine without an
sm_100 device every compute call returns WF_ERR_DEVICE, surfaced as RuntimeError.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

ABI_VERSION = 2  # WF_ABI_VERSION of include/wfb200.h
WF_F32, WF_F64 = 0, 1
ENC_SOQPSK_DIFF, ENC_SOQPSK, ENC_BINARY, ENC_QUATERNARY, ENC_SOQPSK_TERNARY = range(5)

_vp, _i32, _i64, _u64, _dbl = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_double



class LinkCfg(C.Structure):
    """struct wf_link_cfg of include/wfb200.h (field order and types exactly as declared there)."""
    _fields_ = [("mask", C.c_uint64), ("init_state", C.c_uint64), ("period", C.c_uint64), ("degree", C.c_int32),
                ("encoder", C.c_int32), ("n_bits", C.c_int64),
                ("mod_index", C.c_void_p), ("pulse", C.c_void_p), ("n_h", C.c_int32), ("K", C.c_int32),
                ("sps", C.c_int32), ("tune", C.c_int32),
                ("initial_phase", C.c_double), ("derotate", C.c_double), ("seed", C.c_uint64),
                ("taps_re", C.c_void_p), ("n_filt", C.c_int32), ("n_taps", C.c_int32), ("center", C.c_int32),
                ("diff", C.c_int32), ("first_instant", C.c_int64), ("n_calls", C.c_int64),
                ("combine_re", C.c_void_p), ("combine_im", C.c_void_p)]


# name -> (restype, argtypes); order and meaning exactly as in include/wfb200.h
PROTOTYPES = {
    "wf_abi_version": (_i32, []),
    "wf_last_error_string": (C.c_char_p, []),
    "wf_device_check": (_i32, []),
    "wf_lfsr_mask": (_u64, [_i32]),
    "wf_lfsr_jump": (_u64, [_u64, _i32, _u64, _u64]),
    "wf_prbs_encode": (_i32, [_u64, _i32, _u64, _u64, _vp, _u64, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _vp, _vp]),
    "wf_bits_encode": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _i64, _vp, _vp]),
    "wf_trellis_encode": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "wf_cpm_modulate": (_i32, [_vp, _i64, _i64, _vp, _i32, _vp, _i32, _i32, _dbl, _dbl, _i32, _vp, _i64,
                               _vp, _vp, _u64, _u64, _vp, _vp]),
    "wf_cpm_modulate_bank_supported": (_i32, [_i32, _i32, _i32, _i32, _i32, _i64, _i64, _i64]),
    "wf_cpm_modulate_bank": (_i32, [_vp, _i64, _i64, _vp, _i32, _vp, _i32, _i32, _dbl, _dbl, _vp, _vp, _u64, _u64, _vp,
                                    _vp, _i32, _i32, _i32, _i64, _i64, _vp, _i64, _i32, _vp]),
    "wf_frequency_modulate": (_i32, [_vp, _i64, _i64, _i64, _i32, _dbl, _i32, _vp, _i64, _vp]),
    "wf_phase_modulate": (_i32, [_vp, _i64, _dbl, _i32, _vp, _vp]),
    "wf_awgn_add": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _vp, _u64, _u64, _vp, _vp]),
    "wf_awgn_fill": (_i32, [_vp, _i64, _i64, _vp, _vp, _u64, _u64, _vp, _vp]),
    "wf_mf_bank": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _i32, _i32, _i64, _i64,
                          _vp, _i64, _vp]),
    "wf_viterbi_soqpsk": (_i32, [_vp, _i32, _i64, _i64, _i64, _i32, _i32, _vp, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp,
                                 _vp]),
    "wf_viterbi_soqpsk_combined": (_i32, [_vp, _i32, _i64, _i64, _i64, _i32, _vp, _vp, _i32, _vp, _i64, _i64, _vp, _vp,
                                          _vp]),
    "wf_viterbi_cpm": (_i32, [_vp, _i32, _i64, _i64, _i64, _i32, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _i64, _i64, _i64,
                              _vp, _vp, _vp]),
    "wf_cpm_detect_supported": (_i32, [_i32, _i64, _i64, _vp, _vp, _i32, _i64, _i64, _i32, _i32, _i32, _vp, _i32, _i32]),
    "wf_cpm_detect": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _vp, _i32, _i64, _i64, _i32, _i32, _i32, _vp, _i32, _i32,
                             _vp, _vp, _i64, _i64, _i64, _vp, _vp, _i32, _vp]),
    "wf_welch_psd": (_i32, [_vp, _i64, _i64, _i64, _i32, _dbl, _vp, _vp]),
    "wf_eye_traces": (_i32, [_vp, _i64, _i64, _i64, _i32, _vp, _vp]),
    "wf_phase_tree": (_i32, [_vp, _i64, _i64, _i64, _i32, _i32, _dbl, _vp, _vp]),
    "wf_timing_metric": (_i32, [_vp, _i64, _i64, _i64, _i32, _vp, _vp]),
    "wf_bert": (_i32, [_vp, _vp, _i64, _vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "wf_link_sim_workspace_bytes": (_i64, [_vp, _i64]),
    "wf_link_sim_fused": (_i32, [_vp, _i64, _u64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp]),
    "wf_static_scratch_max_calls": (_i64, []),
    "wf_soqpsk_detect_static": (_i32, [_vp, _i64, _i64, _i64, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _i32, _i64, _i64,
                                       _i32, _i32, _vp, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp]),
}


def bind(cdll: C.CDLL) -> C.CDLL:
    """Attach the prototypes to a loaded library; AttributeError if a symbol is missing."""
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(cdll, name)
        fn.restype = res
        fn.argtypes = args
    return cdll


_lock = threading.Lock()
_lib = None


def lib_path() -> str:
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_lib", "libwfb200.so")


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                path = lib_path()
                if not os.path.exists(path):
                    from . import build as _build

                    _build.build()
                _lib = bind(C.CDLL(path))
                if _lib.wf_abi_version() != ABI_VERSION:
                    raise RuntimeError("libwfb200 ABI version mismatch")
    return _lib


class WfError(RuntimeError):
    pass


def check(rc: int) -> None:
    """Raise for a negative wf_status.  WF_ERR_ARG -> ValueError, others -> WfError."""
    if rc == 0:
        return
    msg = lib().wf_last_error_string().decode("utf-8", "replace")
    if rc == -1:
        raise ValueError(msg)
    raise WfError(f"libwfb200 error {rc}: {msg}")


def dbl_array(values):
    """Host double[] argument for the small tables (pulses, taps, mod indices)."""
    vals = [float(v) for v in values]
    return (C.c_double * len(vals))(*vals)
