"""torch-CPU autograd twin of oracle/sketch_oracle.py (fp64 by default).

TEST INFRASTRUCTURE ONLY.  It restates the same reference graph (model.py:34-182) with
torch ops so that autograd provides ground-truth gradients for the manual BPTT in the
NumPy oracle.  parity unpinned (see sketch_oracle.py header).
"""
import math
import numpy as np
import torch

from . import sketch_oracle as so


def _cell(cfg, P, l, inp, state):
    H = cfg.rnn_size
    names = so._cell_names(cfg, l)
    if cfg.model == "lstm":  # BasicLSTMCell, gate order i, j, f, o; forget_bias 1.0
        c, h = state[:, :H], state[:, H:]
        pre = torch.cat([inp, h], 1) @ P[names[0]] + P[names[1]]
        i, j, f, o = pre[:, :H], pre[:, H:2 * H], pre[:, 2 * H:3 * H], pre[:, 3 * H:]
        c2 = c * torch.sigmoid(f + 1.0) + torch.sigmoid(i) * torch.tanh(j)
        h2 = torch.tanh(c2) * torch.sigmoid(o)
        return h2, torch.cat([c2, h2], 1)
    if cfg.model == "gru":  # GRUCell: r,u gates; candidate on [x, r*h]
        h = state
        ru = torch.sigmoid(torch.cat([inp, h], 1) @ P[names[0]] + P[names[1]])
        r, u = ru[:, :H], ru[:, H:]
        cand = torch.tanh(torch.cat([inp, r * h], 1) @ P[names[2]] + P[names[3]])
        h2 = u * h + (1 - u) * cand
        return h2, h2
    h2 = torch.tanh(torch.cat([inp, state], 1) @ P[names[0]] + P[names[1]])
    return h2, h2


def loss_and_grads(cfg, params, x, y, state_in=None, dropout_mask=None, dtype=torch.float64):
    """model.py:66-100 unroll, :142-163 coef, :112-139 loss, :182 tf.gradients."""
    P = {n: torch.tensor(np.asarray(v), dtype=dtype, requires_grad=True) for n, v in params.items()}
    xt = torch.tensor(x, dtype=dtype)
    yt = torch.tensor(y, dtype=dtype)
    B, T, _ = x.shape
    L, M = cfg.num_layers, cfg.num_mixture
    if state_in is None:
        state_in = so.zero_state(cfg, B)
    s0 = [torch.tensor(s, dtype=dtype) for s in state_in]
    state = list(s0)
    outs = []
    for t in range(T):
        inp = xt[:, t, :]
        new = []
        for l in range(L):
            inp, ns = _cell(cfg, P, l, inp, state[l])
            new.append(ns)
        if dropout_mask is not None:
            inp = inp * torch.tensor(dropout_mask[:, t, :], dtype=dtype) / cfg.keep_prob
        outs.append(inp)
        eoc = (xt[:, t, 3] > 0)[:, None]
        state = [torch.where(eoc, s0[l], new[l]) for l in range(L)]
    output = torch.cat(outs, 1).reshape(-1, cfg.rnn_size)
    z = output @ P["rnn_mdn/output_w"] + P["rnn_mdn/output_b"]
    z_pen = z[:, 0:3]
    z_pi, mu1, mu2, s1, s2, rho = torch.split(z[:, 3:], M, dim=1)
    z_pi = torch.exp(z_pi - z_pi.max(1, keepdim=True)[0])
    pi = z_pi * (1.0 / z_pi.sum(1, keepdim=True))
    s1, s2, rho = torch.exp(s1), torch.exp(s2), torch.tanh(rho)
    tgt = yt.reshape(-1, 5)
    x1, x2, pen = tgt[:, 0:1], tgt[:, 1:2], tgt[:, 2:5]
    n1, n2 = x1 - mu1, x2 - mu2
    s1s2 = s1 * s2
    zz = (n1 / s1) ** 2 + (n2 / s2) ** 2 - 2 * (rho * (n1 * n2)) / s1s2
    neg = 1 - rho ** 2
    dens = torch.exp(-zz / (2 * neg)) / (2 * math.pi * s1s2 * torch.sqrt(neg))
    r1 = (dens * pi).sum(1, keepdim=True)
    eps = torch.tensor(1e-20, dtype=dtype)
    cost_shape = (-torch.log(torch.maximum(r1, eps))).mean()
    xent = -(pen * torch.log_softmax(z_pen, 1)).sum(1)
    sif = cfg.stroke_importance_factor
    w = pen[:, 2] + math.sqrt(sif) * pen[:, 0] + sif * pen[:, 1]
    cost_pen = (xent * w).mean()
    cost = cost_shape + cost_pen
    cost.backward()
    grads = {n: (p.grad.numpy().copy() if p.grad is not None else np.zeros(p.shape)) for n, p in P.items()}
    mdn = dict(pi=pi, mu1=mu1, mu2=mu2, sigma1=s1, sigma2=s2, corr=rho, pen=z_pen)
    mdn = {k: v.detach().numpy() for k, v in mdn.items()}
    return ((float(cost), float(cost_shape), float(cost_pen)), grads,
            [s.detach().numpy() for s in state], mdn)
