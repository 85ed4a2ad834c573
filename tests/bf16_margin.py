# bf16 path error levels against the fp64 oracle (margin to the 2e-2 tolerance): python tests/bf16_margin.py
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import sketch_oracle as so
from sketch_rnn_b200 import _lib

def f64(p): return {k: v.astype(np.float64) for k, v in p.items()}
for (H, L, B, T) in [(256, 2, 64, 40), (512, 3, 140, 30), (512, 2, 460, 12)]:
    cfg = so.Config(model="lstm", rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=0.8)
    params = so.init_params(cfg, seed=1, trained_like=True)
    x, y = so.synthetic_batch(B, T, seed=2, p_eoc=0.05)
    s0 = so.synthetic_state(cfg, B, seed=0)
    mask = so.synthetic_dropout_mask(cfg, B, T, seed=0)
    lr, gr, fs, _ = so.loss_and_grads(cfg, f64(params), x.astype(np.float64), y.astype(np.float64),
                                      [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    h = _lib.Handle(model="lstm", rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=0.8,
                    precision="bf16")
    h.set_params_dict(params)
    losses, st = h.loss_grad(x, y, np.ascontiguousarray(np.stack(s0, 0), dtype=np.float32), dropout_mask=mask)
    losses = losses.cpu().numpy()
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    gscale = max(np.max(np.abs(v)) for v in gr.values())
    worst = max(np.max(np.abs(g[n].astype(np.float64) - gr[n])) / gscale for n in gr)
    print("H%d L%d B%d T%d: loss rel err %.2e, worst grad err / max|grad| %.2e" % (H, L, B, T, abs(losses[0] - lr[0]) / abs(lr[0]), worst))
    h.close()
