"""ctypes wrapper of oracle/libdau_cpu.so (TEST INFRASTRUCTURE ONLY, see dau_cpu.c)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libdau_cpu.so")
_fp = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def load():
    if not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE])
    lib = C.CDLL(_SO)
    lib.dau_cpu_num_threads.restype = C.c_int
    lib.dau_cpu_set_num_threads.argtypes = [C.c_int]
    lib.dau_cpu_set_num_threads.restype = None
    lib.dau_cpu_filters.argtypes = [C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int),
                                    np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")]
    lib.dau_cpu_forward.argtypes = [_fp, _fp, _fp, _fp, C.c_float] + [C.c_int] * 12 + [_fp]
    lib.dau_cpu_backward.argtypes = [_fp, _fp, _fp, _fp, _fp, C.c_float] + [C.c_int] * 14 + [_fp] * 5
    return lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def filters(sigma, unit_norm=True, square_norm=False, single_dim=False, forbid_pos=False):
    lib = load()
    kb = C.c_int(0)
    buf = np.zeros(6 * 17 * 17, np.float64)
    assert lib.dau_cpu_filters(sigma, int(unit_norm), int(square_norm), int(single_dim), int(forbid_pos),
                               C.byref(kb), buf) == 0
    k = kb.value
    return [buf[i * k * k:(i + 1) * k * k].reshape(k, k).copy() for i in range(6)]


def forward(x, w, mu1, mu2, sigma, units_ignore=0, interp=True, unit_norm=True, square_norm=False,
            single_dim=False, forbid_pos=False):
    lib = load()
    x, w, mu1, mu2 = map(_f32, (x, w, mu1, mu2))
    N, S, H, W = x.shape
    _, _, G, F = w.shape
    y = np.empty((N, F, H, W), np.float32)
    rc = lib.dau_cpu_forward(x, w, mu1, mu2, float(sigma), N, S, F, G, H, W, units_ignore, int(interp),
                             int(unit_norm), int(square_norm), int(single_dim), int(forbid_pos), y)
    assert rc == 0, rc
    return y


def backward(x, err, w, mu1, mu2, sigma, units_ignore=0, interp=True, unit_norm=True, square_norm=False,
             single_dim=False, forbid_pos=False, mask_last_row=False, mask_last_col=False):
    lib = load()
    x, err, w, mu1, mu2 = map(_f32, (x, err, w, mu1, mu2))
    N, S, H, W = x.shape
    _, _, G, F = w.shape
    dx = np.empty_like(x)
    g = [np.empty_like(w) for _ in range(4)]
    rc = lib.dau_cpu_backward(x, err, w, mu1, mu2, float(sigma), N, S, F, G, H, W, units_ignore, int(interp),
                              int(unit_norm), int(square_norm), int(single_dim), int(forbid_pos),
                              int(mask_last_row), int(mask_last_col), dx, *g)
    assert rc == 0, rc
    return (dx,) + tuple(g)
