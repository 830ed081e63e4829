"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/scan_ref.c (built by oracle/Makefile)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_scan.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.rrl_oracle_segment.restype = ctypes.c_int64
    return _lib


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t)) if a is not None else None


def finish_paths(rew, val, offsets, last_vals, gamma, lam, with_baseline, adv=None, ret=None):
    rew = np.ascontiguousarray(rew, np.float32)
    n = len(rew)
    val = np.ascontiguousarray(val, np.float32) if val is not None else None
    offsets = np.ascontiguousarray(offsets, np.int64)
    last_vals = np.ascontiguousarray(last_vals, np.float32) if last_vals is not None else None
    adv = np.zeros(n, np.float32) if adv is None else adv
    ret = np.zeros(n, np.float32) if ret is None else ret
    lib().rrl_oracle_finish_paths(
        _p(rew, ctypes.c_float), _p(val, ctypes.c_float), _p(offsets, ctypes.c_int64),
        _p(last_vals, ctypes.c_float), ctypes.c_int64(len(offsets) - 1), ctypes.c_double(gamma),
        ctypes.c_double(lam), ctypes.c_int(int(with_baseline)), _p(adv, ctypes.c_float),
        _p(ret, ctypes.c_float))
    return adv, ret


def segment(done, rew, start_ptr=0):
    done = np.ascontiguousarray(done, np.uint8)
    rew = np.ascontiguousarray(rew, np.float32)
    n = len(done)
    offsets = np.zeros(n + 1, np.int64)
    last_vals = np.zeros(max(n, 1), np.float32)
    keep = np.zeros(max(n, 1), np.int64)
    nk = ctypes.c_int64(0)
    ne = lib().rrl_oracle_segment(_p(done, ctypes.c_uint8), _p(rew, ctypes.c_float),
                                  ctypes.c_int64(n), ctypes.c_int64(start_ptr),
                                  _p(offsets, ctypes.c_int64), _p(last_vals, ctypes.c_float),
                                  _p(keep, ctypes.c_int64), ctypes.byref(nk))
    return keep[:nk.value], offsets[:ne + 1], last_vals[:ne]


def normalise(adv):
    adv = np.ascontiguousarray(adv, np.float32)
    n = len(adv)
    tmp = np.empty(n, np.float32)
    out = np.empty(n, np.float32)
    m, s = ctypes.c_float(0), ctypes.c_float(0)
    lib().rrl_oracle_normalise(_p(adv, ctypes.c_float), ctypes.c_int64(n), _p(tmp, ctypes.c_float),
                               _p(out, ctypes.c_float), ctypes.byref(m), ctypes.byref(s))
    return out, np.float32(m.value), np.float32(s.value)
