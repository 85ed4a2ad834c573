"""ctypes loader for oracle/sampler_ref.c (TEST INFRASTRUCTURE ONLY, see that file's header)."""
import ctypes as C
import os
import subprocess

import numpy as np

from . import sketch_oracle as so

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libskref.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "_build/libskref.so"])
    return _LIB


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "sampler_ref.c")
        if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
            build()
        _lib = C.CDLL(_LIB)
        _lib.skref_sample.restype = C.c_int
        for f in ("skref_exp", "skref_log", "skref_tanh"):
            getattr(_lib, f).restype = C.c_float
            getattr(_lib, f).argtypes = [C.c_float]
    return _lib


def _offsets(cfg):
    """Flat-vector offsets per layer in oracle param_spec order."""
    off, o = {}, 0
    for n, s in so.param_spec(cfg):
        off[n] = o
        o += int(np.prod(s))
    w0, b0, w1, b1 = [], [], [], []
    for l in range(cfg.num_layers):
        names = so._cell_names(cfg, l)
        w0.append(off[names[0]])
        b0.append(off[names[1]])
        w1.append(off[names[2]] if len(names) > 2 else 0)
        b1.append(off[names[3]] if len(names) > 2 else 0)
    return w0, b0, w1, b1, off["rnn_mdn/output_w"], off["rnn_mdn/output_b"]


def sample(cfg, params, uniforms, num=300, temp_mixture=1.0, temp_pen=1.0, stop_if_eoc=False,
           want_params=False):
    """Model.sample for one sketch (reference model.py:187-264) through the C restatement.
    uniforms: float32 [num, 4].  Returns (strokes[n,5], mixp[n,NOUT] or None)."""
    L = lib()
    flat = np.ascontiguousarray(so.flatten(cfg, params), dtype=np.float32)
    uniforms = np.ascontiguousarray(uniforms, dtype=np.float32)
    assert uniforms.shape == (num, 4)
    w0, b0, w1, b1, ow, ob = _offsets(cfg)
    arr = lambda v: (C.c_int64 * len(v))(*v)
    strokes = np.zeros((num, 5), dtype=np.float32)
    mixp = np.zeros((num, cfg.nout), dtype=np.float32) if want_params else None
    cell = {"lstm": 0, "gru": 1, "rnn": 2}[cfg.model]
    n = L.skref_sample(C.c_int(cell), C.c_int(cfg.rnn_size), C.c_int(cfg.num_layers),
                       C.c_int(cfg.num_mixture), flat.ctypes.data_as(C.c_void_p), arr(w0), arr(b0),
                       arr(w1), arr(b1), C.c_int64(ow), C.c_int64(ob), C.c_int(num),
                       C.c_float(temp_mixture), C.c_float(temp_pen), C.c_int(1 if stop_if_eoc else 0),
                       C.c_float(cfg.data_scale), uniforms.ctypes.data_as(C.c_void_p),
                       strokes.ctypes.data_as(C.c_void_p),
                       mixp.ctypes.data_as(C.c_void_p) if want_params else None)
    return strokes[:n].copy(), (mixp[:n].copy() if want_params else None)
