"""Regenerates tests/golden/*.npz: known-answer vectors of the oracle (fp64 maths, stored as
float32/float64) for small configs of every cell type.  The reference itself holds no golden
vectors for this path and cannot run here (Python 2 + TensorFlow 0.x), so these pin the oracle
against regressions and give the CUDA path fixed inputs/outputs to be compared with; they were
produced by `python tests/golden/make_golden.py` in this repository.

Each file holds: the config, flat parameters (TF variable order), x, y, state_in, dropout mask,
and the oracle's losses, flat gradient, final state, MDN parameters, plus one sampler trace from the
C restatement (uniforms -> strokes)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sketch_oracle as so  # noqa: E402
from oracle import sampler_ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [("lstm", 16, 2, 3, 4, 6), ("gru", 16, 2, 3, 4, 6), ("rnn", 16, 2, 3, 4, 6), ("lstm", 24, 3, 5, 3, 5)]


def make(model, H, L, M, B, T, seed=0):
    cfg = so.Config(model=model, rnn_size=H, num_layers=L, num_mixture=M, batch_size=B, seq_length=T,
                    keep_prob=0.8)
    params = so.init_params(cfg, seed=seed + 1, trained_like=True)
    x, y = so.synthetic_batch(B, T, seed=seed + 2, p_eoc=0.2)
    s0 = so.synthetic_state(cfg, B, seed=seed)
    mask = so.synthetic_dropout_mask(cfg, B, T, seed=seed)
    p64 = {k: v.astype(np.float64) for k, v in params.items()}
    losses, grads, fstate, mdn = so.loss_and_grads(cfg, p64, x.astype(np.float64), y.astype(np.float64),
                                                   [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    fw = so.forward(cfg, p64, x.astype(np.float64), [s.astype(np.float64) for s in s0])
    u = np.random.RandomState(seed + 3).rand(24, 4).astype(np.float32)
    strokes, mixp = sampler_ref.sample(cfg, params, u, 24, 0.4, 0.4, False, want_params=True)
    names = ["pi", "mu1", "mu2", "sigma1", "sigma2", "corr"]
    out = dict(model=model, H=H, L=L, M=M, B=B, T=T, params=so.flatten(cfg, params), x=x, y=y,
               state_in=np.stack(s0, 0), mask=mask, losses=np.array(losses, dtype=np.float64),
               grads=so.flatten(cfg, grads, dtype=np.float64), final_state=np.stack(fstate, 0),
               mdn_nodrop=np.concatenate([fw["mdn"]["pen"]] + [fw["mdn"][n] for n in names], axis=1),
               final_state_nodrop=np.stack(fw["final_state"], 0),
               sample_uniforms=u, sample_strokes=strokes, sample_mixp=mixp)
    np.savez_compressed(os.path.join(HERE, "%s_h%d_l%d.npz" % (model, H, L)), **out)


if __name__ == "__main__":
    for c in CASES:
        make(*c)
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
