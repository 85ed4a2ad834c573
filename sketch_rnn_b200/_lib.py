"""ctypes binding of libskrnn.so (include/skrnn.h).

There is no CPU fallback: if the shared library is missing, or no B200 is visible, the
calls below raise.  PyTorch is used only to own device memory and streams.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libskrnn.so")

CELL = {"lstm": 0, "gru": 1, "rnn": 2}
PREC = {"f32": 0, "bf16": 1}


class SkrnnError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("cell", C.c_int32), ("rnn_size", C.c_int32), ("num_layers", C.c_int32),
                ("num_mixture", C.c_int32), ("batch_size", C.c_int32), ("seq_length", C.c_int32),
                ("keep_prob", C.c_float), ("grad_clip", C.c_float),
                ("stroke_importance_factor", C.c_float), ("data_scale", C.c_float),
                ("precision", C.c_int32), ("device", C.c_int32)]


# every symbol include/skrnn.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "skrnn_last_error": (C.c_char_p, []),
    "skrnn_version": (C.c_char_p, []),
    "skrnn_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "skrnn_destroy": (None, [_P]),
    "skrnn_param_count": (C.c_int64, [_P]),
    "skrnn_num_tensors": (C.c_int32, [_P]),
    "skrnn_tensor_info": (C.c_int, [_P, C.c_int32, C.c_char_p, C.c_int32, C.POINTER(C.c_int64),
                                    C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "skrnn_set_params": (C.c_int, [_P, _P, C.c_int64]),
    "skrnn_get_params": (C.c_int, [_P, _P, C.c_int64]),
    "skrnn_set_adam_state": (C.c_int, [_P, _P, _P, C.c_int64, C.c_double, C.c_double]),
    "skrnn_get_adam_state": (C.c_int, [_P, _P, _P, C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "skrnn_params_dev": (_P, [_P]),
    "skrnn_grads_dev": (_P, [_P]),
    "skrnn_forward": (C.c_int, [_P, C.c_int32, C.c_int32, _P, _P, _P, _P, _P]),
    "skrnn_loss_grad": (C.c_int, [_P, C.c_int32, C.c_int32, _P, _P, _P, _P, C.c_uint64, C.c_int32,
                                  C.c_int64, _P, _P, _P]),
    "skrnn_grad_segment": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "skrnn_stream_wait_grad": (C.c_int, [_P, C.c_int32, _P]),
    "skrnn_apply_clip_adam": (C.c_int, [_P, C.c_float, _P, _P]),
    "skrnn_train_step_host": (C.c_int, [_P, C.c_int32, C.c_int32, _P, _P, _P, C.c_float, C.c_uint64, _P]),
    "skrnn_sample": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32, _P, C.c_uint64,
                               C.c_int32, _P, _P, _P, _P]),
    "skrnn_launch_count": (C.c_int64, [_P]),
    "skrnn_clip_adam_noalloc": (C.c_int, [C.c_int64, _P, _P, _P, _P, C.c_float, C.c_float, C.c_double, C.c_double,
                                          _P, _P]),
    "skrnn_crc32c": (C.c_uint32, [C.c_uint32, C.c_char_p, C.c_size_t]),
    "skrnn_debug_gemm_bf16": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, _P, C.c_int32, C.c_int32, _P, C.c_int32,
                                        C.c_int32, _P, C.c_int32, _P, C.c_int32, _P, _P]),
    "skrnn_profile_enable": (C.c_int, [_P, C.c_int32]),
    "skrnn_profile_read": (C.c_int, [_P, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
}

_lib = None


def load():
    """dlopen libskrnn.so and declare prototypes.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SkrnnError("%s is missing: run `python -m sketch_rnn_b200.build` "
                         "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise SkrnnError(load().skrnn_last_error().decode("utf-8", "replace"))


def _np_ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Handle(object):
    """Owns one skrnn_handle (weights, optimiser state, workspaces on one GPU)."""

    def __init__(self, model="lstm", rnn_size=256, num_layers=2, num_mixture=24, batch_size=0,
                 seq_length=0, keep_prob=1.0, grad_clip=5.0, stroke_importance_factor=200.0,
                 data_scale=15.0, precision="f32", device=0):
        if model not in CELL:
            raise Exception("model type not supported: {}".format(model))  # reference model.py:21
        self.lib = load()
        self.cfg = Config(CELL[model], rnn_size, num_layers, num_mixture, batch_size, seq_length,
                          keep_prob, grad_clip, stroke_importance_factor, data_scale,
                          PREC[precision], device)
        self.model = model
        self.H, self.L, self.M = rnn_size, num_layers, num_mixture
        self.NOUT = 3 + 6 * num_mixture
        self.S = 2 * rnn_size if model == "lstm" else rnn_size
        self.device = device
        self._h = _P()
        check(self.lib.skrnn_create(C.byref(self.cfg), C.byref(self._h)))
        self.P = int(self.lib.skrnn_param_count(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.skrnn_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- parameters ---------------------------------------------------------------------
    def tensors(self):
        """[(tf_name, offset, shape)] of the flat parameter vector."""
        out = []
        buf = C.create_string_buffer(256)
        off, r, c = C.c_int64(), C.c_int32(), C.c_int32()
        for i in range(self.lib.skrnn_num_tensors(self._h)):
            check(self.lib.skrnn_tensor_info(self._h, i, buf, 256, C.byref(off), C.byref(r), C.byref(c)))
            name = buf.value.decode()
            shape = (c.value,) if name.endswith(("/Bias", "output_b")) else (r.value, c.value)
            out.append((name, off.value, shape))
        return out

    def set_params(self, flat):
        flat = np.ascontiguousarray(flat, dtype=np.float32)
        check(self.lib.skrnn_set_params(self._h, _np_ptr(flat), flat.size))

    def get_params(self):
        flat = np.empty(self.P, dtype=np.float32)
        check(self.lib.skrnn_get_params(self._h, _np_ptr(flat), flat.size))
        return flat

    def set_params_dict(self, params):
        flat = np.empty(self.P, dtype=np.float32)
        for name, off, shape in self.tensors():
            a = np.asarray(params[name], dtype=np.float32)
            assert a.shape == tuple(shape), (name, a.shape, shape)
            flat[off:off + a.size] = a.ravel()
        self.set_params(flat)

    def get_params_dict(self, flat=None):
        flat = self.get_params() if flat is None else flat
        return {n: flat[o:o + int(np.prod(s))].reshape(s).copy() for n, o, s in self.tensors()}

    def set_adam_state(self, m, v, beta1_power, beta2_power):
        m = np.ascontiguousarray(m, dtype=np.float32)
        v = np.ascontiguousarray(v, dtype=np.float32)
        check(self.lib.skrnn_set_adam_state(self._h, _np_ptr(m), _np_ptr(v), m.size, beta1_power, beta2_power))

    def get_adam_state(self):
        m = np.empty(self.P, dtype=np.float32)
        v = np.empty(self.P, dtype=np.float32)
        b1, b2 = C.c_double(), C.c_double()
        check(self.lib.skrnn_get_adam_state(self._h, _np_ptr(m), _np_ptr(v), m.size, C.byref(b1), C.byref(b2)))
        return m, v, b1.value, b2.value

    def grads_tensor(self):
        """torch view of the handle-owned flat gradient (for an external all-reduce)."""
        return _wrap_device_ptr(self.lib.skrnn_grads_dev(self._h), self.P, self.device)

    def params_tensor(self):
        return _wrap_device_ptr(self.lib.skrnn_params_dev(self._h), self.P, self.device)

    # -- compute (torch CUDA tensors in/out) -------------------------------------------------
    def forward(self, x, state_in, want_mdn=True, stream=None):
        import torch
        B, T = int(x.shape[0]), int(x.shape[1])
        x = _dev(x, self.device)
        state_in = _dev(state_in, self.device)
        assert tuple(state_in.shape) == (self.L, B, self.S), state_in.shape
        mdn = torch.empty((B * T, self.NOUT), dtype=torch.float32, device=x.device) if want_mdn else None
        state_out = torch.empty_like(state_in)
        check(self.lib.skrnn_forward(self._h, B, T, x.data_ptr(), state_in.data_ptr(),
                                     mdn.data_ptr() if want_mdn else None, state_out.data_ptr(),
                                     _stream_ptr(stream)))
        return mdn, state_out

    def loss_grad(self, x, y, state_in, dropout_mask=None, seed=0, row0=0, loss_rows=None, stream=None):
        import torch
        B, T = int(x.shape[0]), int(x.shape[1])
        x, y, state_in = _dev(x, self.device), _dev(y, self.device), _dev(state_in, self.device)
        mask = _dev(dropout_mask, self.device) if dropout_mask is not None else None
        losses = torch.empty(3, dtype=torch.float32, device=x.device)
        state_out = torch.empty_like(state_in)
        check(self.lib.skrnn_loss_grad(self._h, B, T, x.data_ptr(), y.data_ptr(), state_in.data_ptr(),
                                       mask.data_ptr() if mask is not None else None, seed, row0,
                                       B * T if loss_rows is None else loss_rows, losses.data_ptr(),
                                       state_out.data_ptr(), _stream_ptr(stream)))
        return losses, state_out

    def grad_segments(self):
        """[(index, offset, count)] in the order skrnn_loss_grad finishes them: output layer (index L), then layers
        L-1 .. 0."""
        out = []
        off, cnt = C.c_int64(), C.c_int64()
        for idx in [self.L] + list(range(self.L - 1, -1, -1)):
            check(self.lib.skrnn_grad_segment(self._h, idx, C.byref(off), C.byref(cnt)))
            out.append((idx, off.value, cnt.value))
        return out

    def stream_wait_grad(self, index, stream):
        """`stream` (torch.cuda.Stream) waits until the last loss_grad has finished gradient segment `index`."""
        check(self.lib.skrnn_stream_wait_grad(self._h, index, C.c_void_p(stream.cuda_stream)))

    def apply_clip_adam(self, lr, stream=None):
        import torch
        gn = torch.empty(1, dtype=torch.float32, device="cuda:%d" % self.device)
        check(self.lib.skrnn_apply_clip_adam(self._h, lr, gn.data_ptr(), _stream_ptr(stream)))
        return gn

    def train_step_host(self, x, y, state, lr, seed=0):
        """One sess.run([cost, cost_shape, cost_pen, final_state, train_op]) with numpy buffers.
        `state` ([L,B,S] float32) is updated in place; returns the three losses."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.ascontiguousarray(y, dtype=np.float32)
        assert state.dtype == np.float32 and state.flags["C_CONTIGUOUS"]
        B, T = x.shape[0], x.shape[1]
        losses = np.empty(3, dtype=np.float32)
        check(self.lib.skrnn_train_step_host(self._h, B, T, _np_ptr(x), _np_ptr(y), _np_ptr(state),
                                             lr, seed, _np_ptr(losses)))
        return losses

    def sample(self, n, steps, temp_mixture=1.0, temp_pen=1.0, stop_if_eoc=False, uniforms=None,
               seed=0, sketch0=0, want_params=False, stream=None):
        import torch
        dev = "cuda:%d" % self.device
        strokes = torch.empty((n, steps, 5), dtype=torch.float32, device=dev)
        lengths = torch.empty((n,), dtype=torch.int32, device=dev)
        mixp = torch.empty((n, steps, self.NOUT), dtype=torch.float32, device=dev) if want_params else None
        if uniforms is not None:
            uniforms = _dev(uniforms, self.device)
            assert tuple(uniforms.shape) == (n, steps, 4), uniforms.shape
        check(self.lib.skrnn_sample(self._h, n, steps, temp_mixture, temp_pen, 1 if stop_if_eoc else 0,
                                    uniforms.data_ptr() if uniforms is not None else None, seed, sketch0,
                                    strokes.data_ptr(), lengths.data_ptr(),
                                    mixp.data_ptr() if want_params else None, _stream_ptr(stream)))
        return strokes, lengths, mixp

    def debug_gemm_bf16(self, a, a_kmajor, b, b_kmajor, M, N, K, bias=None, accumulate=False, c=None,
                        overrides=None):
        """C[M,N] (+)= A*B through the tcgen05 GEMM.  a, b: 2-D torch bf16 CUDA tensors in their
        storage layout ([M,K] or [K,M] resp. [N,K] or [K,N])."""
        import torch
        assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
        if c is None:
            c = torch.zeros((M, N), dtype=torch.float32, device=a.device)
        ov = (C.c_uint32 * 6)(*overrides) if overrides is not None else None
        check(self.lib.skrnn_debug_gemm_bf16(self._h, M, N, K, a.data_ptr(), a.stride(0), 1 if a_kmajor else 0,
                                             b.data_ptr(), b.stride(0), 1 if b_kmajor else 0, c.data_ptr(),
                                             c.stride(0), bias.data_ptr() if bias is not None else None,
                                             1 if accumulate else 0, ov, _stream_ptr(None)))
        return c

    def launch_count(self):
        return int(self.lib.skrnn_launch_count(self._h))

    PROFILE_CATEGORIES = ("gemm_f32", "rnn_fwd", "rnn_bwd", "mdn_head", "clip_adam", "misc", "sampler", "gemm_tc")

    def profile_enable(self, on=True):
        check(self.lib.skrnn_profile_enable(self._h, 1 if on else 0))

    def profile_read(self):
        """{category: (total_ms, launches)} since the last read (synchronises the device)."""
        ms = (C.c_double * 8)()
        cnt = (C.c_int64 * 8)()
        check(self.lib.skrnn_profile_read(self._h, ms, cnt))
        return {n: (ms[i], int(cnt[i])) for i, n in enumerate(self.PROFILE_CATEGORIES)}


NOALLOC_MAX_PARAMS = 8192 * 4096


def clip_adam_noalloc(grads, params, m, v, lr, grad_clip, beta1_power, beta2_power, gnorm_out=None, stream=None):
    """Clip + Adam on caller-owned torch CUDA float32 tensors, updated in place, without any device
    allocation inside the library (skrnn_clip_adam_noalloc).  For callers whose tensors come from a
    framework pool and who cannot create a Handle (that would cudaMalloc).  At most NOALLOC_MAX_PARAMS
    elements; calls on one device must not overlap (shared fixed scratch)."""
    n = grads.numel()
    for t in (grads, params, m, v):
        assert t.is_cuda and t.is_contiguous() and t.numel() == n and str(t.dtype) == "torch.float32"
    check(load().skrnn_clip_adam_noalloc(n, grads.data_ptr(), params.data_ptr(), m.data_ptr(), v.data_ptr(),
                                         lr, grad_clip, beta1_power, beta2_power,
                                         gnorm_out.data_ptr() if gnorm_out is not None else None,
                                         _stream_ptr(stream)))


def _dev(a, device):
    import torch
    if isinstance(a, np.ndarray):
        a = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    a = a.to(device="cuda:%d" % device, dtype=torch.float32)
    return a.contiguous()


def _stream_ptr(stream):
    import torch
    if stream is None:
        stream = torch.cuda.current_stream()
    return C.c_void_p(stream.cuda_stream)


def _wrap_device_ptr(ptr, n, device):
    """Zero-copy torch float32 view of n floats at a raw device pointer."""
    import torch

    class _Arr(object):
        pass

    a = _Arr()
    a.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(a, device="cuda:%d" % device)
