"""`Model(args, infer)` and a `Session` stand-in with the reference's Python surface (reference model.py,
train.py:57-86, sample.py:57-64), on top of libskrnn.so.

The reference builds a TensorFlow graph and evaluates pieces of it with `sess.run(fetches, feed)`.
Here every graph attribute the callers use (`model.cost`, `model.final_state`, `model.train_op`, ...) is a
small token and `Session.run` maps the reference's four call patterns onto the C ABI:

    sess.run(model.assign_lr(v))                                 train.py:78   (tf.assign(model.lr, v))
    sess.run(model.initial_state)                                train.py:81
    sess.run([model.cost, model.cost_shape, model.cost_pen,      train.py:86
              model.final_state, model.train_op], feed)
    sess.run([model.pi, ..., model.pen, model.final_state], feed) model.py:215-217

There is no CPU fallback; constructing a Model needs a B200 and the built library.
"""
import math
import os
import pickle

import numpy as np

from . import _lib
from . import utils


class Fetch(object):
    """A named node of the reference graph."""

    def __init__(self, name):
        self.name = name

    def __repr__(self):
        return "<skrnn fetch %s>" % self.name


class _AssignLr(object):
    def __init__(self, value):
        self.value = float(value)


class Model(object):
    def __init__(self, args, infer=False, precision=None, device=0, seed=0):
        if infer:  # reference model.py:9-11 mutates the caller's args
            args.batch_size = 1
            args.seq_length = 1
        self.args = args
        if args.model not in ("rnn", "gru", "lstm"):
            raise Exception("model type not supported: {}".format(args.model))  # model.py:21
        self.infer = infer
        self.num_mixture = args.num_mixture
        self.NOUT = 3 + self.num_mixture * 6  # model.py:43
        precision = precision or getattr(args, "precision", None) or os.environ.get("SKRNN_PRECISION", "bf16")
        self.precision = "f32" if infer else precision  # the sampler runs the bit-exact fp32 path
        # training-mode dropout only (model.py:29-30)
        keep = args.keep_prob if (not infer and args.keep_prob < 1) else 1.0
        self.handle = _lib.Handle(model=args.model, rnn_size=args.rnn_size, num_layers=args.num_layers,
                                  num_mixture=args.num_mixture, batch_size=args.batch_size,
                                  seq_length=args.seq_length, keep_prob=keep, grad_clip=args.grad_clip,
                                  stroke_importance_factor=args.stroke_importance_factor,
                                  data_scale=args.data_scale, precision=self.precision, device=device)
        self.state_size = self.handle.S
        # graph attributes the reference callers touch
        self.input_data = Fetch("input_data")
        self.target_data = Fetch("target_data")
        self.initial_state = Fetch("initial_state")
        self.final_state = Fetch("final_state")
        self.cost = Fetch("cost")
        self.cost_shape = Fetch("cost_shape")
        self.cost_pen = Fetch("cost_pen")
        self.train_op = Fetch("train_op")
        for n in ("pi", "mu1", "mu2", "sigma1", "sigma2", "corr", "pen"):
            setattr(self, n, Fetch(n))
        self.cell = self  # `self.cell.zero_state(...)` (model.py:208)
        self.initial_input = np.zeros((args.batch_size, 5), dtype=np.float32)
        self.initial_input[:, 4] = 1.0  # built but never used by the reference either (model.py:52-54)
        self.lr = 0.01  # tf.Variable(0.01, trainable=False), model.py:180
        self._step = 0
        self.seed = seed
        self.initialize_all_variables(seed)

    # -- variables -------------------------------------------------------------------------------
    def initialize_all_variables(self, seed=0):
        """tf.initialize_all_variables(): TF 0.x default initialiser (uniform_unit_scaling, +-sqrt(3/fan_in));
        _linear biases start at zero except the GRU gate bias (1.0); Adam slots zero."""
        flat = utils.default_init(self.handle, seed)
        self.handle.set_params(flat)
        self.handle.set_adam_state(np.zeros_like(flat), np.zeros_like(flat), 0.9, 0.999)
        self.lr = 0.01
        self._step = 0

    def assign_lr(self, value):
        """Stands in for tf.assign(model.lr, value) (train.py:78)."""
        return _AssignLr(value)

    def zero_state(self, batch_size, dtype=None):
        """cell.zero_state (model.py:36,208): tuple of L arrays [B, S]."""
        return tuple(np.zeros((batch_size, self.state_size), dtype=np.float32) for _ in range(self.args.num_layers))

    # -- the two evaluations behind Session.run --------------------------------------------------------
    def _stack_state(self, state, B):
        if state is None:
            state = self.zero_state(B)
        st = np.ascontiguousarray(np.stack([np.asarray(s, dtype=np.float32) for s in state], 0))
        assert st.shape == (self.args.num_layers, B, self.state_size), st.shape
        return st

    def _train_state(self, state, B):
        """The [L, B, S] state buffer of a training step.  train.py feeds back the final_state tuple of the previous
        step (train.py:84-86): that tuple's arrays are views of this model's page-locked buffer, which the step updates in
        place, so nothing is stacked or copied.  Any other state is copied in.  (A caller that keeps a returned
        final_state across later steps must copy it: the views follow the buffer.)"""
        shape = (self.args.num_layers, B, self.state_size)
        buf = getattr(self, "_st_buf", None)
        if buf is None or buf.shape != shape:
            try:
                import torch
                buf = torch.empty(shape, dtype=torch.float32).pin_memory().numpy()
            except Exception:
                buf = np.empty(shape, dtype=np.float32)
            self._st_buf, self._st_tuple = buf, None
        if state is None or state is not self._st_tuple:
            buf[...] = self._stack_state(state, B)
        return buf

    def train_step(self, x, y, state=None):
        """One sess.run([cost, cost_shape, cost_pen, final_state, train_op], feed) (train.py:85-86)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        st = self._train_state(state, x.shape[0])
        self._step += 1
        losses = self.handle.train_step_host(x, y, st, self.lr, seed=(self.seed << 32) + self._step)
        self._st_tuple = tuple(st[l] for l in range(st.shape[0]))
        return float(losses[0]), float(losses[1]), float(losses[2]), self._st_tuple

    def forward(self, x, state=None):
        """sess.run([pi, mu1, mu2, sigma1, sigma2, corr, pen, final_state], feed) (model.py:215-217)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        B = x.shape[0]
        st = self._stack_state(state, B)
        mdn, state_out = self.handle.forward(x, st)
        mdn = mdn.cpu().numpy()
        M = self.num_mixture
        out = {"pen": mdn[:, 0:3]}
        for i, n in enumerate(("pi", "mu1", "mu2", "sigma1", "sigma2", "corr")):
            out[n] = mdn[:, 3 + i * M:3 + (i + 1) * M]
        so = state_out.cpu().numpy()
        out["final_state"] = tuple(so[l] for l in range(so.shape[0]))
        return out

    # -- sampler (model.py:187-264) ----------------------------------------------------------------------
    def sample(self, sess=None, num=300, temp_mixture=1.0, temp_pen=1.0, stop_if_eoc=False, uniforms=None):
        """Same signature and return value as the reference: (strokes [n,5] float32 with dx,dy already
        multiplied by data_scale, mixture_params = per point [pi_pdf, mu1, mu2, sigma1, sigma2, corr, pen_pdf]).
        The whole loop runs in one kernel launch; the four uniforms per point come from np.random
        (the reference draws random.random() x2 and np.random.multivariate_normal)."""
        if uniforms is None:
            uniforms = np.random.rand(1, num, 4).astype(np.float32)
        uniforms = np.ascontiguousarray(uniforms, dtype=np.float32).reshape(1, num, 4)
        strokes, lengths, mixp = self.handle.sample(1, num, temp_mixture, temp_pen, stop_if_eoc, uniforms=uniforms,
                                                    want_params=True)
        n = int(lengths.cpu().numpy()[0])
        strokes = strokes.cpu().numpy()[0, :n].copy()
        mixp = mixp.cpu().numpy()[0, :n]
        return strokes, self.mixture_params(mixp)

    def sample_batch(self, n_sketches, num=300, temp_mixture=1.0, temp_pen=1.0, stop_if_eoc=False, seed=0,
                     uniforms=None, sketch0=0):
        """Many sketches in one launch.  Returns (strokes [n, num, 5] with zero rows past each length, lengths [n])."""
        strokes, lengths, _ = self.handle.sample(n_sketches, num, temp_mixture, temp_pen, stop_if_eoc,
                                                 uniforms=uniforms, seed=seed, sketch0=sketch0)
        return strokes.cpu().numpy(), lengths.cpu().numpy()

    def mixture_params(self, mixp):
        """[n, NOUT] rows as the kernel returns them -> the reference's per-point list
        [pi_pdf, mu1, mu2, sigma1, sigma2, corr, pen_pdf] (model.py:251-252)."""
        M = self.num_mixture
        return [[mixp[i, 3:3 + M], mixp[i, 3 + M:3 + 2 * M], mixp[i, 3 + 2 * M:3 + 3 * M], mixp[i, 3 + 3 * M:3 + 4 * M],
                 mixp[i, 3 + 4 * M:3 + 5 * M], mixp[i, 3 + 5 * M:3 + 6 * M], mixp[i, 0:3]] for i in range(len(mixp))]

    def sample_candidates(self, n_sketches, num=300, temp_mixture=1.0, temp_pen=1.0, stop_if_eoc=False, uniforms=None):
        """`n_sketches` independent Model.sample calls in ONE launch: returns (strokes [n, num, 5] with zero rows past
        each length, lengths [n], mixp [n, num, NOUT]) as NumPy arrays.  Candidate k consumes the uniforms
        np.random.rand(1, num, 4) that the k-th sequential `sample` call would have drawn (one rand(n, num, 4) call
        yields the same numbers in the same order), so a loop over `sample` and one call here draw identical sketches."""
        if uniforms is None:
            uniforms = np.random.rand(n_sketches, num, 4).astype(np.float32)
        uniforms = np.ascontiguousarray(uniforms, dtype=np.float32).reshape(n_sketches, num, 4)
        strokes, lengths, mixp = self.handle.sample(n_sketches, num, temp_mixture, temp_pen, stop_if_eoc,
                                                    uniforms=uniforms, want_params=True)
        return strokes.cpu().numpy(), lengths.cpu().numpy(), mixp.cpu().numpy()

    # -- checkpoints: save/<dataset>/{checkpoint, model.ckpt-N} (train.py:72-75, sample.py:59-64) --------
    def variables(self):
        """tf.all_variables() by TF name: trainables, Adam slots `<var>/Adam`, `<var>/Adam_1`, the
        beta powers and the learning-rate variable `Variable`."""
        out = {}
        p = self.handle.get_params_dict()
        m, v, b1, b2 = self.handle.get_adam_state()
        md, vd = self.handle.get_params_dict(m), self.handle.get_params_dict(v)
        for n in p:
            out[n] = p[n]
            out[n + "/Adam"] = md[n]
            out[n + "/Adam_1"] = vd[n]
        out["beta1_power"] = np.float32(b1)
        out["beta2_power"] = np.float32(b2)
        out["Variable"] = np.float32(self.lr)
        # not a TF variable: the number of training steps taken, which seeds the in-kernel dropout mask of the next
        # step (a restored model must not replay the masks it has already used)
        out["skrnn/dropout_step"] = np.float32(self._step)
        return out

    def load_variables(self, var):
        names = [n for n, _, _ in self.handle.tensors()]
        self.handle.set_params_dict({n: var[n] for n in names})
        if all((n + "/Adam") in var for n in names):  # slots are optional when only sampling
            flat_m = np.zeros(self.handle.P, dtype=np.float32)
            flat_v = np.zeros(self.handle.P, dtype=np.float32)
            for n, off, shape in self.handle.tensors():
                k = int(np.prod(shape))
                flat_m[off:off + k] = np.asarray(var[n + "/Adam"], dtype=np.float32).ravel()
                flat_v[off:off + k] = np.asarray(var[n + "/Adam_1"], dtype=np.float32).ravel()
            self.handle.set_adam_state(flat_m, flat_v, float(var.get("beta1_power", 0.9)),
                                       float(var.get("beta2_power", 0.999)))
        if "Variable" in var:
            self.lr = float(var["Variable"])
        if "skrnn/dropout_step" in var:
            self._step = int(var["skrnn/dropout_step"])


class Session(object):
    """Just enough of tf.Session / tf.InteractiveSession for train.py and sample.py."""

    def __init__(self, model=None):
        self.model = model

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def run(self, fetches, feed_dict=None):
        m = self.model
        if isinstance(fetches, _AssignLr):
            m.lr = fetches.value
            return m.lr
        single = not isinstance(fetches, (list, tuple))
        flist = [fetches] if single else list(fetches)
        feed = {}
        for k, v in (feed_dict or {}).items():
            feed[k.name if isinstance(k, Fetch) else k] = v
        names = [f.name for f in flist]
        if names == ["initial_state"] and not feed:
            res = {"initial_state": m.zero_state(m.args.batch_size)}
        elif "train_op" in names:
            c, cs, cp, st = m.train_step(feed["input_data"], feed["target_data"], feed.get("initial_state"))
            res = {"cost": c, "cost_shape": cs, "cost_pen": cp, "final_state": st, "train_op": None}
        else:
            res = m.forward(feed["input_data"], feed.get("initial_state"))
        out = [res[n] for n in names]
        return out[0] if single else out


class Saver(object):
    """tf.train.Saver over `save/<dataset>/`: writes `model.ckpt-<step>` and the text index file `checkpoint`
    (model_checkpoint_path / all_model_checkpoint_paths lines, as TF does), keeping the newest `max_to_keep`
    checkpoints of this Saver and deleting older ones (TF's default is 5)."""

    def __init__(self, model, max_to_keep=5):
        self.model = model
        self.max_to_keep = max_to_keep
        self._kept = []

    def save(self, sess, checkpoint_path, global_step=0):
        from . import checkpoint as ck
        path = "%s-%d" % (checkpoint_path, global_step)
        ck.write_checkpoint(path, self.model.variables())
        if path in self._kept:
            self._kept.remove(path)
        self._kept.append(path)
        while self.max_to_keep and len(self._kept) > self.max_to_keep:
            old = self._kept.pop(0)
            try:
                os.remove(old)
            except OSError:
                pass
        d = os.path.dirname(path)
        with open(os.path.join(d, "checkpoint"), "w") as f:
            f.write('model_checkpoint_path: "%s"\n' % os.path.basename(path))
            for k in self._kept:
                f.write('all_model_checkpoint_paths: "%s"\n' % os.path.basename(k))
        return path

    def restore(self, sess, path):
        from . import checkpoint as ck
        self.model.load_variables(ck.read_checkpoint(path))


def get_checkpoint_state(dirname):
    """tf.train.get_checkpoint_state: parses `<dir>/checkpoint`; returns an object with
    .model_checkpoint_path or None."""
    idx = os.path.join(dirname, "checkpoint")
    if not os.path.exists(idx):
        return None
    path = None
    for line in open(idx):
        if line.startswith("model_checkpoint_path:"):
            path = line.split(":", 1)[1].strip().strip('"')
    if path is None:
        return None
    if not os.path.isabs(path):
        path = os.path.join(dirname, path)

    class _State(object):
        pass

    s = _State()
    s.model_checkpoint_path = path
    return s


def load_config(path):
    """config.pkl written by train.py:54-55 (Python 2 text-mode protocol-0 pickle of the argparse Namespace)."""
    with open(path, "rb") as f:
        return pickle.load(f, encoding="latin1")


def save_config(path, args):
    with open(path, "wb") as f:
        pickle.dump(args, f, protocol=2)
