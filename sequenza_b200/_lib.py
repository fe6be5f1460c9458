"""ctypes binding of libsequenza_b200.so (include/sequenza_b200.h).

There is no CPU implementation behind this module: if the shared object is missing the
import of any compute function raises, and every entry point returns SQZ_ERR_NOGPU
(raised as RuntimeError) when no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libsequenza_b200.so")

SQZ_MODE_FAST = 0
SQZ_MODE_FAITHFUL = 1

_d, _i, _i64, _p, _sz = C.c_double, C.c_int, C.c_int64, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); mirrors include/sequenza_b200.h declaration by declaration
SIGNATURES = {
    "sqz_version": (C.c_char_p, []),
    "sqz_last_error": (C.c_char_p, []),
    "sqz_device_count": (_i, []),
    "sqz_release": (_i, []),
    "sqz_baf_grid": (_i, [_p] * 6 + [_i64, _p, _i64, _p, _i64, _d, _i, _i, _i, _p, _p, _i,
                          _i, _i, _i, _p]),
    "sqz_baf_grid_batch": (_i, [_i64, _p] + [_p] * 6 + [_p, _i64, _p, _i64, _p, _i, _i, _i,
                                _p, _p, _i, _i, _i, _i, _p]),
    "sqz_baf_grid_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64]),
    "sqz_baf_grid_workspace_preferred_bytes": (_sz, [_i64, _i64, _i64, _i64, _i64, _i, _sz]),
    "sqz_baf_grid_units_workspace_bytes": (_sz, [_i64, _i64, _p, _i64, _i64, _i, _i, _i, _i,
                                                 _i64, _i64, _sz]),
    "sqz_baf_grid_dev_units": (_i, [_i64, _p, _i64] + [_p] * 6 + [_p, _i64, _p, _i64, _i64,
                                    _i64, _p, _i, _i, _i, _p, _p, _i, _i, _i, _p, _sz, _p, _p,
                                    _i]),
    "sqz_baf_grid_dev": (_i, [_i64, _p, _i64] + [_p] * 6 + [_p, _i64, _p, _i64, _i64, _i64,
                              _p, _i, _i, _i, _p, _p, _i, _i, _i, _p, _sz, _p, _p]),
    "sqz_baf_grid_dev_noalloc": (_i, [_i64, _p, _i64] + [_p] * 6 + [_p, _i64, _p, _i64, _i64,
                                      _i64, _p, _i, _i, _i, _p, _p, _i, _i, _i, _p, _p]),
    "sqz_peer_data_offset": (_sz, []),
    "sqz_peer_alloc": (_i, [_sz, _p, _p, _p]),
    "sqz_peer_open": (_i, [_p, _p]),
    "sqz_peer_close": (_i, [_p]),
    "sqz_peer_free": (_i, [_p]),
    "sqz_peer_signal": (_i, [_p, _i, _i, C.c_uint64, _p]),
    "sqz_peer_wait": (_i, [_p, _i, C.c_uint64, _p]),
    "sqz_baf_point": (_i, [_p] * 6 + [_i64, _d, _d, _d, _i, _i, _i, _p, _p, _i, _i, _p]),
    "sqz_baf_point_dev": (_i, [_p] * 6 + [_i64, _d, _d, _d, _i, _i, _i, _p, _p, _i, _i, _p,
                               _p]),
    "sqz_mufreq_point": (_i, [_p] * 4 + [_i64, _d, _d, _d, _i, _i, _i, _p, _p, _i, _p]),
    "sqz_mufreq_scalar_table_bytes": (_sz, [_d, _d]),
    "sqz_mufreq_point_scalar": (_i, [_p, _p, _i64, _d, _d, _d, _d, _d, _i, _i, _i, _p, _p, _i,
                                     _p]),
    "sqz_mufreq_point_scalar_dev": (_i, [_p, _p, _i64, _d, _d, _d, _d, _d, _i, _i, _i, _p, _p,
                                         _i, _p, _sz, _p, _p]),
    "sqz_mufreq_point_dev": (_i, [_p] * 4 + [_i64, _d, _d, _d, _i, _i, _i, _p, _p, _i, _p,
                                  _p]),
    "sqz_mufreq_grid": (_i, [_p] * 4 + [_i64, _p, _i64, _p, _i64, _d, _i, _i, _i, _p, _p,
                             _i, _i, _p]),
    "sqz_mufreq_grid_dev": (_i, [_p] * 4 + [_i64, _p, _i64, _p, _i64, _i64, _i64, _d, _i,
                                 _i, _i, _p, _p, _i, _p, _sz, _p, _p]),
    "sqz_mufreq_grid_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "sqz_fit_finish": (_i, [_p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p]),
    "sqz_fit_finish_dev": (_i, [_p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "sqz_segment_stats": (_i, [_i64, _p, _p] + [_p] * 8 + [_d, _i, _p]),
    "sqz_segment_stats_workspace_bytes": (_sz, [_i64, _i64]),
    "sqz_segment_stats_dev": (_i, [_i64, _p, _p, _i64, _i64] + [_p] * 8 + [_d, _i, _p, _sz, _p,
                                   _p]),
    "sqz_model_points": (_i, [_i, _d, _d, _d, _i, _i, _i, _p, _p, _i, _p, _p]),
    "sqz_fp64_peak": (_i, [_i, _i, _p]),
    "sqz_baf_grid_plan": (_i, [_i64, _i64, _p, _i64, _i64, _i, _i, _i, _i, _p]),
    "sqz_profile_enable": (_i, [_i]),
    "sqz_profile_collect": (_i, [_p, _p]),
    "sqz_profile_counters": (_i, [_p, _p, _p]),
}

_lib = None


def lib():
    """The loaded library.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(
                "libsequenza_b200.so is missing (%s). Build it with "
                "`python -m sequenza_b200.build`; there is no CPU fallback." % SO_PATH)
        handle = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int):
    if rc != 0:
        msg = lib().sqz_last_error().decode("utf8", "replace")
        if rc == -1:
            raise ValueError(msg)  # R: stop()
        raise RuntimeError("libsequenza_b200 error %d: %s" % (rc, msg))


def vec(x, n=None, name="argument"):
    """as.double + R's scalar recycling into a contiguous float64 vector."""
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1))
    if n is not None and a.size != n:
        if a.size == 1:
            a = np.full(n, a[0], dtype=np.float64)
        else:
            raise ValueError("%s has length %d, expected 1 or %d" % (name, a.size, n))
    return a


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None
