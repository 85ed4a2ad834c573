"""CPU oracle for sketch-rnn's training step and sampling loop (NumPy).

TEST INFRASTRUCTURE ONLY.  Nothing in the shipped package imports this file;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs do.

This is a restatement of the graph that ``/root/reference/model.py`` builds with
TensorFlow 0.x (which is not vendored in the reference and cannot run here), written
from the reference call sites and from the published TF r0.10-0.12 semantics of the
ops it calls.  Every function cites the reference lines it follows.

PARITY STATUS: **parity unpinned** -- the reference has no tests, golden vectors or
fixtures for this path and its pretrained ``save/kanji/model.ckpt-0`` blob is absent
(SURVEY.md section 8c).  What pins this oracle instead:
  * ``oracle/sketch_oracle_torch.py``: the same graph under torch fp64 autograd, which
    validates the manual backward pass here to ~1e-10;
  * central finite differences on a tiny config (tests/test_oracle.py);
  * committed known-answer vectors under ``tests/golden/`` made by
    ``tests/golden/make_golden.py``.

Conventions:
  x, y           [B, T, 5] float  = [dx, dy, eos, eoc, cont]      (model.py:34-35)
  state          list of L arrays [B, S]; S = 2H for lstm ([c | h], TF BasicLSTMCell
                 state_is_tuple=False), H for gru / rnn                (model.py:36)
  params         dict name -> array, names follow TF variable scopes (see param_spec)
  head rows      index b*T + t                                        (model.py:98)
"""
import math
import numpy as np

NG = {"lstm": 4, "gru": 3, "rnn": 1}


class Config(object):
    """Hyper-parameters read by Model.__init__ (model.py:8-47, 134, 182)."""

    def __init__(self, model="lstm", rnn_size=256, num_layers=2, num_mixture=24,
                 batch_size=100, seq_length=300, keep_prob=0.8, grad_clip=5.0,
                 stroke_importance_factor=200.0, data_scale=15.0):
        if model not in NG:
            # model.py:21
            raise Exception("model type not supported: {}".format(model))
        self.model = model
        self.rnn_size = int(rnn_size)
        self.num_layers = int(num_layers)
        self.num_mixture = int(num_mixture)
        self.batch_size = int(batch_size)
        self.seq_length = int(seq_length)
        self.keep_prob = float(keep_prob)
        self.grad_clip = float(grad_clip)
        self.stroke_importance_factor = float(stroke_importance_factor)
        self.data_scale = float(data_scale)

    @property
    def nout(self):
        return 3 + 6 * self.num_mixture  # model.py:43

    @property
    def state_size(self):
        return 2 * self.rnn_size if self.model == "lstm" else self.rnn_size


# ----------------------------------------------------------------------------------
# parameters
# ----------------------------------------------------------------------------------
def param_spec(cfg):
    """[(tf_variable_name, shape)] in TF creation order: output_w/output_b first
    (model.py:45-47), then one Cell{l} scope per layer (model.py:27,77; TF MultiRNNCell).
    Matrix rows are [x rows ; h rows] (TF _linear on concat([inputs, state]))."""
    H, L = cfg.rnn_size, cfg.num_layers
    spec = [("rnn_mdn/output_w", (H, cfg.nout)), ("rnn_mdn/output_b", (cfg.nout,))]
    for l in range(L):
        I = 5 if l == 0 else H
        base = "rnn_mdn/MultiRNNCell/Cell%d/" % l
        if cfg.model == "lstm":
            spec += [(base + "BasicLSTMCell/Linear/Matrix", (I + H, 4 * H)),
                     (base + "BasicLSTMCell/Linear/Bias", (4 * H,))]
        elif cfg.model == "gru":
            spec += [(base + "GRUCell/Gates/Linear/Matrix", (I + H, 2 * H)),
                     (base + "GRUCell/Gates/Linear/Bias", (2 * H,)),
                     (base + "GRUCell/Candidate/Linear/Matrix", (I + H, H)),
                     (base + "GRUCell/Candidate/Linear/Bias", (H,))]
        else:
            spec += [(base + "BasicRNNCell/Linear/Matrix", (I + H, H)),
                     (base + "BasicRNNCell/Linear/Bias", (H,))]
    return spec


def param_count(cfg):
    return int(sum(int(np.prod(s)) for _, s in param_spec(cfg)))


def init_params(cfg, seed=0, dtype=np.float32, trained_like=False):
    """TF-0.x default initialiser (uniform_unit_scaling: +-sqrt(3/fan_in), fan_in =
    prod(shape[:-1])), biases from _linear are zero except GRU gate bias = 1.0.
    trained_like=True rescales the head so sigma-hat ~ [-2,0] and |rho-hat| < 1.5, which
    exercises the clamp / underflow branches."""
    rng = np.random.RandomState(seed)
    out = {}
    for name, shape in param_spec(cfg):
        if name.endswith("/Bias"):
            v = np.zeros(shape)
            if "Gates" in name:
                v += 1.0
        else:
            fan_in = float(np.prod(shape[:-1])) if len(shape) > 1 else 1.0
            lim = math.sqrt(3.0 / fan_in)
            v = rng.uniform(-lim, lim, size=shape)
        out[name] = v.astype(dtype)
    if trained_like:
        M = cfg.num_mixture
        b = out["rnn_mdn/output_b"].astype(np.float64)
        b[3 + 3 * M:3 + 5 * M] = rng.uniform(-2.0, 0.0, size=2 * M)
        b[3 + 5 * M:3 + 6 * M] = rng.uniform(-1.5, 1.5, size=M)
        out["rnn_mdn/output_b"] = b.astype(dtype)
        out["rnn_mdn/output_w"] = (out["rnn_mdn/output_w"] * 2.0).astype(dtype)
    return out


def flatten(cfg, params, dtype=np.float32):
    return np.concatenate([np.asarray(params[n], dtype=dtype).ravel() for n, _ in param_spec(cfg)])


def unflatten(cfg, flat):
    out, o = {}, 0
    for n, s in param_spec(cfg):
        k = int(np.prod(s))
        out[n] = np.array(flat[o:o + k]).reshape(s)
        o += k
    assert o == flat.size
    return out


def _cell_names(cfg, l):
    base = "rnn_mdn/MultiRNNCell/Cell%d/" % l
    if cfg.model == "lstm":
        return (base + "BasicLSTMCell/Linear/Matrix", base + "BasicLSTMCell/Linear/Bias")
    if cfg.model == "gru":
        return (base + "GRUCell/Gates/Linear/Matrix", base + "GRUCell/Gates/Linear/Bias",
                base + "GRUCell/Candidate/Linear/Matrix", base + "GRUCell/Candidate/Linear/Bias")
    return (base + "BasicRNNCell/Linear/Matrix", base + "BasicRNNCell/Linear/Bias")


def zero_state(cfg, batch, dtype=np.float32):
    """cell.zero_state (model.py:36)."""
    return [np.zeros((batch, cfg.state_size), dtype=dtype) for _ in range(cfg.num_layers)]


# ----------------------------------------------------------------------------------
# cells (TF r0.10-0.12 rnn_cell.py semantics; call site model.py:77)
# ----------------------------------------------------------------------------------
def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def _cell_fwd(cfg, params, l, inp, state):
    """One cell call. Returns (out, new_state, cache)."""
    H = cfg.rnn_size
    names = _cell_names(cfg, l)
    if cfg.model == "lstm":
        W, b = params[names[0]], params[names[1]]
        c, h = state[:, :H], state[:, H:]
        xh = np.concatenate([inp, h], axis=1)
        pre = xh.dot(W) + b
        i = _sigmoid(pre[:, 0:H])
        j = np.tanh(pre[:, H:2 * H])
        f = _sigmoid(pre[:, 2 * H:3 * H] + 1.0)  # forget_bias=1.0 added at run time
        o = _sigmoid(pre[:, 3 * H:4 * H])
        c2 = c * f + i * j
        tc = np.tanh(c2)
        h2 = tc * o
        return h2, np.concatenate([c2, h2], axis=1), (xh, c, i, j, f, o, tc)
    if cfg.model == "gru":
        Wg, bg, Wc, bc = (params[n] for n in names)
        h = state
        xh = np.concatenate([inp, h], axis=1)
        ru = _sigmoid(xh.dot(Wg) + bg)
        r, u = ru[:, :H], ru[:, H:]
        xrh = np.concatenate([inp, r * h], axis=1)
        cand = np.tanh(xrh.dot(Wc) + bc)
        h2 = u * h + (1.0 - u) * cand
        return h2, h2, (xh, xrh, h, r, u, cand)
    W, b = params[names[0]], params[names[1]]
    h = state
    xh = np.concatenate([inp, h], axis=1)
    h2 = np.tanh(xh.dot(W) + b)
    return h2, h2, (xh, h2)


def _cell_bwd(cfg, params, grads, l, cache, dout, dstate):
    """Reverse of _cell_fwd. dout: grad wrt cell output, dstate: grad wrt new_state.
    Accumulates weight grads into `grads`; returns (dinp, dstate_prev)."""
    H = cfg.rnn_size
    names = _cell_names(cfg, l)
    if cfg.model == "lstm":
        W = params[names[0]]
        xh, c, i, j, f, o, tc = cache
        dc2 = dstate[:, :H]
        dh2 = dstate[:, H:] + dout
        do = dh2 * tc
        dc2 = dc2 + dh2 * o * (1.0 - tc * tc)
        di, dj, df, dc = dc2 * j, dc2 * i, dc2 * c, dc2 * f
        dpre = np.concatenate([di * i * (1 - i), dj * (1 - j * j), df * f * (1 - f), do * o * (1 - o)], axis=1)
        grads[names[0]] += xh.T.dot(dpre)
        grads[names[1]] += dpre.sum(0)
        dxh = dpre.dot(W.T)
        I = xh.shape[1] - H
        return dxh[:, :I], np.concatenate([dc, dxh[:, I:]], axis=1)
    if cfg.model == "gru":
        Wg, Wc = params[names[0]], params[names[2]]
        xh, xrh, h, r, u, cand = cache
        I = xh.shape[1] - H
        dh2 = dstate + dout
        du = dh2 * (h - cand)
        dcand = dh2 * (1.0 - u)
        dh = dh2 * u
        dprec = dcand * (1 - cand * cand)
        grads[names[2]] += xrh.T.dot(dprec)
        grads[names[3]] += dprec.sum(0)
        dxrh = dprec.dot(Wc.T)
        dinp = dxrh[:, :I].copy()
        drh = dxrh[:, I:]
        dr = drh * h
        dh = dh + drh * r
        dpreg = np.concatenate([dr * r * (1 - r), du * u * (1 - u)], axis=1)
        grads[names[0]] += xh.T.dot(dpreg)
        grads[names[1]] += dpreg.sum(0)
        dxh = dpreg.dot(Wg.T)
        dinp += dxh[:, :I]
        dh = dh + dxh[:, I:]
        return dinp, dh
    W = params[names[0]]
    xh, h2 = cache
    I = xh.shape[1] - H
    dpre = (dstate + dout) * (1 - h2 * h2)
    grads[names[0]] += xh.T.dot(dpre)
    grads[names[1]] += dpre.sum(0)
    dxh = dpre.dot(W.T)
    return dxh[:, :I], dxh[:, I:]


# ----------------------------------------------------------------------------------
# unroll with end-of-content state reset (model.py:66-95)
# ----------------------------------------------------------------------------------
def unroll_fwd(cfg, params, x, state_in, dropout_mask=None):
    """custom_rnn_autodecoder.  Per step: run the whole layer stack, take the top layer's
    (pre-reset) output, apply DropoutWrapper's output dropout (model.py:29-30: only on the
    stack output, never on state), then for every layer state <- where(x_t[:,3] > 0,
    FED initial_state, new_state) (model.py:82-92; generalised from [0,1] to range(L))."""
    B, T, _ = x.shape
    L, H = cfg.num_layers, cfg.rnn_size
    dt = x.dtype
    state = [s.astype(dt) for s in state_in]
    outs = np.zeros((B, T, H), dtype=dt)
    caches = []
    for t in range(T):
        inp = x[:, t, :]
        step_cache = []
        new_states = []
        for l in range(L):
            out, ns, cache = _cell_fwd(cfg, params, l, inp, state[l])
            step_cache.append(cache)
            new_states.append(ns)
            inp = out
        if dropout_mask is not None:
            inp = inp * dropout_mask[:, t, :].astype(dt) * dt.type(1.0 / cfg.keep_prob)
        outs[:, t, :] = inp
        eoc = x[:, t, 3] > 0
        state = [np.where(eoc[:, None], state_in[l].astype(dt), new_states[l]) for l in range(L)]
        caches.append((step_cache, eoc))
    return outs, state, caches


def unroll_bwd(cfg, params, grads, x, caches, douts, dropout_mask=None):
    """BPTT through unroll_fwd (what tf.gradients does at model.py:182).  final_state gets
    no gradient (it is only fetched, train.py:86); the reset blocks the state gradient where
    the step's input had eoc > 0."""
    B, T, _ = x.shape
    L = cfg.num_layers
    dt = douts.dtype
    dstate = [np.zeros((B, cfg.state_size), dtype=dt) for _ in range(L)]
    for t in reversed(range(T)):
        step_cache, eoc = caches[t]
        keep = (~eoc)[:, None].astype(dt)
        dnew = [d * keep for d in dstate]
        dout = douts[:, t, :]
        if dropout_mask is not None:
            dout = dout * dropout_mask[:, t, :].astype(dt) * dt.type(1.0 / cfg.keep_prob)
        for l in reversed(range(L)):
            dinp, dprev = _cell_bwd(cfg, params, grads, l, step_cache[l], dout, dnew[l])
            dstate[l] = dprev
            dout = dinp
    return


# ----------------------------------------------------------------------------------
# MDN head + loss (model.py:98-178)
# ----------------------------------------------------------------------------------
def mixture_coef(cfg, logits):
    """get_mixture_coef (model.py:142-163).  Returns pi, mu1, mu2, sigma1, sigma2, corr, pen."""
    M = cfg.num_mixture
    z_pen = logits[:, 0:3]
    blocks = [logits[:, 3 + k * M:3 + (k + 1) * M] for k in range(6)]
    z_pi, mu1, mu2, s1h, s2h, rh = blocks
    z_pi = np.exp(z_pi - z_pi.max(axis=1, keepdims=True))
    pi = z_pi * (1.0 / z_pi.sum(axis=1, keepdims=True))
    return pi, mu1, mu2, np.exp(s1h), np.exp(s2h), np.tanh(rh), z_pen


def normal_2d(x1, x2, mu1, mu2, s1, s2, rho):
    """tf_2d_normal (model.py:112-122)."""
    n1 = x1 - mu1
    n2 = x2 - mu2
    s1s2 = s1 * s2
    z = np.square(n1 / s1) + np.square(n2 / s2) - 2 * (rho * (n1 * n2)) / s1s2
    neg = 1 - np.square(rho)
    return np.exp(-z / (2 * neg)) / (2 * np.pi * (s1s2 * np.sqrt(neg)))


def head_fwd(cfg, params, outs, y):
    """xw_plus_b + get_mixture_coef + get_lossfunc (model.py:98-105, 124-139, 165-178)."""
    B, T, H = outs.shape
    dt = outs.dtype
    flat = outs.reshape(B * T, H)
    logits = flat.dot(params["rnn_mdn/output_w"]) + params["rnn_mdn/output_b"]
    pi, mu1, mu2, s1, s2, rho, z_pen = mixture_coef(cfg, logits)
    tgt = y.reshape(B * T, 5)
    x1, x2 = tgt[:, 0:1], tgt[:, 1:2]
    pen_data = tgt[:, 2:5]
    dens = normal_2d(x1, x2, mu1, mu2, s1, s2, rho)
    S = (dens * pi).sum(axis=1, keepdims=True)
    loss_shape_rows = -np.log(np.maximum(S, dt.type(1e-20)))
    cost_shape = loss_shape_rows.mean()
    zmax = z_pen.max(axis=1, keepdims=True)
    lse = zmax + np.log(np.exp(z_pen - zmax).sum(axis=1, keepdims=True))
    logp = z_pen - lse
    xent = -(pen_data * logp).sum(axis=1)
    sif = cfg.stroke_importance_factor
    w = pen_data[:, 2] + dt.type(math.sqrt(sif)) * pen_data[:, 0] + dt.type(sif) * pen_data[:, 1]
    cost_pen = (xent * w).mean()
    cache = (flat, logits, pi, mu1, mu2, s1, s2, rho, dens, S, logp, w, x1, x2, pen_data)
    mdn = dict(pi=pi, mu1=mu1, mu2=mu2, sigma1=s1, sigma2=s2, corr=rho, pen=z_pen, logits=logits)
    return (cost_shape + cost_pen, cost_shape, cost_pen), mdn, cache


def head_bwd(cfg, params, grads, cache, B, T):
    """Analytic backward of head_fwd. Returns d(outs) [B,T,H]."""
    flat, logits, pi, mu1, mu2, s1, s2, rho, dens, S, logp, w, x1, x2, pen_data = cache
    dt = flat.dtype
    N = flat.shape[0]
    M = cfg.num_mixture
    live = (S >= dt.type(1e-20)).astype(dt)  # d max(S,1e-20)/dS: TF routes grad to S when S >= eps
    gamma = live * (pi * dens) / np.maximum(S, dt.type(1e-300) if dt == np.float64 else dt.type(1e-37))
    a = (x1 - mu1) / s1
    b = (x2 - mu2) / s2
    q = 1 - rho * rho
    Z = a * a + b * b - 2 * rho * a * b
    invN = dt.type(1.0 / N)
    dl = np.zeros_like(logits)
    dl[:, 3:3 + M] = (live * pi - gamma) * invN
    dl[:, 3 + M:3 + 2 * M] = -gamma * (a - rho * b) / (s1 * q) * invN
    dl[:, 3 + 2 * M:3 + 3 * M] = -gamma * (b - rho * a) / (s2 * q) * invN
    dl[:, 3 + 3 * M:3 + 4 * M] = -gamma * (a * (a - rho * b) / q - 1) * invN
    dl[:, 3 + 4 * M:3 + 5 * M] = -gamma * (b * (b - rho * a) / q - 1) * invN
    dl[:, 3 + 5 * M:3 + 6 * M] = -gamma * (a * b + rho * (1 - Z / q)) * invN
    # softmax_cross_entropy_with_logits backprop: (softmax * sum(labels) - labels); labels are
    # one-hot in every batch SketchLoader makes, sum(labels) kept for generality.
    p = np.exp(logp)
    dl[:, 0:3] = w[:, None] * (p * pen_data.sum(axis=1, keepdims=True) - pen_data) * invN
    grads["rnn_mdn/output_w"] += flat.T.dot(dl)
    grads["rnn_mdn/output_b"] += dl.sum(0)
    return dl.dot(params["rnn_mdn/output_w"].T).reshape(B, T, -1), dl


# ----------------------------------------------------------------------------------
# the graph: forward / loss / gradients / optimiser
# ----------------------------------------------------------------------------------
def forward(cfg, params, x, state_in=None, y=None, dropout_mask=None):
    """Everything `sess.run` evaluates on the forward side (model.py:97-178)."""
    if state_in is None:
        state_in = zero_state(cfg, x.shape[0], x.dtype)
    outs, final_state, caches = unroll_fwd(cfg, params, x, state_in, dropout_mask)
    if y is None:
        y = np.zeros_like(x)
        y[:, :, 4] = 1
    losses, mdn, hcache = head_fwd(cfg, params, outs, y)
    return dict(losses=losses, mdn=mdn, outs=outs, final_state=final_state,
                _caches=caches, _hcache=hcache)


def loss_and_grads(cfg, params, x, y, state_in=None, dropout_mask=None, loss_scale_rows=None):
    """cost/cost_shape/cost_pen + tf.gradients(cost, tvars) (model.py:176-182), unclipped.
    loss_scale_rows: if given, the means divide by this row count instead of B*T (used to
    emulate one data-parallel shard of a larger global batch)."""
    B, T, _ = x.shape
    fw = forward(cfg, params, x, state_in, y, dropout_mask)
    grads = {n: np.zeros(s, dtype=x.dtype) for n, s in param_spec(cfg)}
    douts, _ = head_bwd(cfg, params, grads, fw["_hcache"], B, T)
    unroll_bwd(cfg, params, grads, x, fw["_caches"], douts, dropout_mask)
    losses = fw["losses"]
    if loss_scale_rows is not None:
        s = x.dtype.type(float(B * T) / float(loss_scale_rows))
        losses = tuple(v * s for v in losses)
        grads = {k: v * s for k, v in grads.items()}
    return losses, grads, fw["final_state"], fw["mdn"]


def global_norm(grads):
    return math.sqrt(sum(float(np.sum(np.square(g.astype(np.float64)))) for g in grads.values()))


def clip_by_global_norm(grads, clip):
    """tf.clip_by_global_norm (model.py:182): g * clip * min(1/gn, 1/clip)."""
    names = list(grads.keys())
    dt = grads[names[0]].dtype
    gn = dt.type(math.sqrt(sum(float(np.sum(np.square(g), dtype=np.float64)) for g in grads.values())))
    scale = dt.type(clip) * min(dt.type(1.0) / gn if gn > 0 else dt.type(np.inf), dt.type(1.0 / clip))
    return {n: g * scale for n, g in grads.items()}, gn


class AdamState(object):
    """Slots of tf.train.AdamOptimizer(lr, epsilon=0.001) (model.py:183)."""

    def __init__(self, cfg, dtype=np.float32):
        self.m = {n: np.zeros(s, dtype=dtype) for n, s in param_spec(cfg)}
        self.v = {n: np.zeros(s, dtype=dtype) for n, s in param_spec(cfg)}
        self.beta1_power = 0.9
        self.beta2_power = 0.999


def adam_apply(params, grads, opt, lr, beta1=0.9, beta2=0.999, eps=1e-3):
    """TF ApplyAdam (training_ops): alpha = lr*sqrt(1-b2^t)/(1-b1^t); m += (g-m)(1-b1);
    v += (g*g-v)(1-b2); var -= alpha*m/(sqrt(v)+eps); then the beta powers advance."""
    for n in params:
        dt = params[n].dtype
        alpha = dt.type(lr * math.sqrt(1.0 - opt.beta2_power) / (1.0 - opt.beta1_power))
        g = grads[n]
        opt.m[n] = opt.m[n] + (g - opt.m[n]) * dt.type(1.0 - beta1)
        opt.v[n] = opt.v[n] + (g * g - opt.v[n]) * dt.type(1.0 - beta2)
        params[n] = params[n] - alpha * opt.m[n] / (np.sqrt(opt.v[n]) + dt.type(eps))
    opt.beta1_power *= beta1
    opt.beta2_power *= beta2


def train_step(cfg, params, opt, x, y, state_in, lr, dropout_mask=None):
    """One sess.run([cost, cost_shape, cost_pen, final_state, train_op]) (train.py:86)."""
    losses, grads, final_state, _ = loss_and_grads(cfg, params, x, y, state_in, dropout_mask)
    clipped, gn = clip_by_global_norm(grads, cfg.grad_clip)
    adam_apply(params, clipped, opt, lr)
    return losses, final_state, gn


# ----------------------------------------------------------------------------------
# sampler (model.py:187-264)
# ----------------------------------------------------------------------------------
def _pi_idx(u, pdf):
    """get_pi_idx (model.py:189-197): first i with cumsum >= u, else -1.  The accumulator is a
    Python/NumPy-1.10 float64 scalar."""
    acc = 0.0
    for i in range(pdf.size):
        acc += float(pdf[i])
        if acc >= u:
            return i
    return -1


def box_muller(u1, u2):
    """Declared deviation from np.random.multivariate_normal (model.py:199-203): the draw is a
    function of two injected uniforms. u1 is mapped to (0,1] so log() is finite."""
    r = np.sqrt(np.float32(-2.0) * np.log(np.float32(1.0) - np.float32(u1)))
    ang = np.float32(2.0 * np.pi) * np.float32(u2)
    return np.float32(r * np.cos(ang)), np.float32(r * np.sin(ang))


def sample(cfg, params, uniforms, num=300, temp_mixture=1.0, temp_pen=1.0, stop_if_eoc=False):
    """Model.sample for one sketch.  uniforms: [num, 4] = (u_pi, u_pen, u_g1, u_g2) replacing
    random.random() x2 (model.py:226,235) and np.random.multivariate_normal (model.py:202).
    Quirks kept: zero first input (model.py:205), temperature only for i > 1 (model.py:220),
    `pi_pdf /= temp_pen` (model.py:230: pen temperature is never applied; the returned pi_pdf
    is scaled), get_pi_idx == -1 -> last component / cont (model.py:240-247), Gaussian not
    temperature-scaled, and the eoc reset holding the state (model.py:82-92 with
    initial_state = prev_state, model.py:215)."""
    f32 = np.float32
    prev_x = np.zeros((1, 1, 5), dtype=f32)
    prev_state = zero_state(cfg, 1, f32)
    strokes = np.zeros((num, 5), dtype=f32)
    mixture_params = []
    n_out = num
    for i in range(num):
        fw = forward(cfg, params, prev_x, prev_state)
        mdn = fw["mdn"]
        next_state = fw["final_state"]
        pi_pdf = mdn["pi"][0].astype(f32).copy()
        if i > 1:
            pi_pdf = np.log(pi_pdf) / f32(temp_mixture)
            pi_pdf -= pi_pdf.max()
            pi_pdf = np.exp(pi_pdf)
        pi_pdf /= pi_pdf.sum()
        idx = _pi_idx(float(uniforms[i, 0]), pi_pdf)
        pen_pdf = mdn["pen"][0].astype(f32).copy()
        if i > 1:
            pi_pdf = pi_pdf / f32(temp_pen)
        pen_pdf -= pen_pdf.max()
        pen_pdf = np.exp(pen_pdf)
        pen_pdf /= pen_pdf.sum()
        pen_idx = _pi_idx(float(uniforms[i, 1]), pen_pdf)
        eos = eoc = cont = 0
        if pen_idx == 0:
            eos = 1
        elif pen_idx == 1:
            eoc = 1
        else:
            cont = 1
        mu1, mu2 = mdn["mu1"][0][idx], mdn["mu2"][0][idx]
        s1, s2, rho = mdn["sigma1"][0][idx], mdn["sigma2"][0][idx], mdn["corr"][0][idx]
        z1, z2 = box_muller(uniforms[i, 2], uniforms[i, 3])
        nx1 = f32(mu1 + s1 * z1)
        nx2 = f32(mu2 + s2 * (rho * z1 + np.sqrt(f32(1.0) - rho * rho) * z2))
        strokes[i, :] = [nx1, nx2, eos, eoc, cont]
        mixture_params.append([pi_pdf, mdn["mu1"][0], mdn["mu2"][0], mdn["sigma1"][0],
                               mdn["sigma2"][0], mdn["corr"][0], pen_pdf])
        if stop_if_eoc and eoc == 1:
            n_out = i + 1
            break
        prev_x = np.zeros((1, 1, 5), dtype=f32)
        prev_x[0, 0] = strokes[i]
        prev_state = next_state
    strokes = strokes[:n_out].copy()
    strokes[:, 0:2] *= f32(cfg.data_scale)
    return strokes, mixture_params


# ----------------------------------------------------------------------------------
# synthetic inputs (the dataset blob is absent)
# ----------------------------------------------------------------------------------
def synthetic_batch(B, T, seed=0, dtype=np.float32, p_eos=0.10, p_eoc=0.015):
    rng = np.random.RandomState(seed)
    seq = np.zeros((B, T + 1, 5), dtype=np.float64)
    seq[:, :, 0:2] = rng.normal(0.0, 0.5, size=(B, T + 1, 2))
    u = rng.rand(B, T + 1)
    pen = np.where(u < p_eoc, 1, np.where(u < p_eoc + p_eos, 0, 2))
    seq[:, :, 2] = pen == 0
    seq[:, :, 3] = pen == 1
    seq[:, :, 4] = pen == 2
    seq = seq.astype(dtype)
    return np.ascontiguousarray(seq[:, :-1]), np.ascontiguousarray(seq[:, 1:])


def synthetic_state(cfg, B, seed=0, dtype=np.float32, scale=0.3):
    rng = np.random.RandomState(seed + 1000)
    return [(rng.uniform(-1, 1, size=(B, cfg.state_size)) * scale).astype(dtype) for _ in range(cfg.num_layers)]


def synthetic_dropout_mask(cfg, B, T, seed=0):
    rng = np.random.RandomState(seed + 2000)
    return (rng.rand(B, T, cfg.rnn_size) < cfg.keep_prob).astype(np.float32)
