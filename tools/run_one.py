# one loss_grad at a given shape (for ncu / SKRNN_TC_DEBUG=1): python tools/run_one.py H L B T [prec [model]]
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sketch_rnn_b200 import _lib, utils
H, L, B, T = [int(v) for v in sys.argv[1:5]]
prec = sys.argv[5] if len(sys.argv) > 5 else "bf16"
model = sys.argv[6] if len(sys.argv) > 6 else "lstm"
h = _lib.Handle(model=model, rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=0.8, precision=prec)
h.set_params(utils.default_init(h, seed=0))
x, y = utils.synthetic_batch(B, T, seed=1)
st = torch.zeros((L, B, h.S), device="cuda")
for i in range(2):
    losses, _ = h.loss_grad(x, y, st, seed=i)
    h.apply_clip_adam(0.005)
torch.cuda.synchronize()
print("loss", losses.cpu().numpy())
