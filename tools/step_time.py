# time of loss_grad + clip_adam at a given shape: python tools/step_time.py H L B T [prec [model [iters]]]
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sketch_rnn_b200 import _lib, utils
H, L, B, T = [int(v) for v in sys.argv[1:5]]
prec = sys.argv[5] if len(sys.argv) > 5 else "bf16"
model = sys.argv[6] if len(sys.argv) > 6 else "lstm"
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 10
h = _lib.Handle(model=model, rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=0.8, precision=prec)
h.set_params(utils.default_init(h, seed=0))
x, y = utils.synthetic_batch(B, T, seed=1)
st = torch.zeros((L, B, h.S), device="cuda")
for i in range(3):
    losses, _ = h.loss_grad(x, y, st, seed=i)
    h.apply_clip_adam(0.005)
torch.cuda.synchronize()
h.profile_enable(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(iters):
    losses, _ = h.loss_grad(x, y, st, seed=10 + i)
    h.apply_clip_adam(0.005)
e1.record()
torch.cuda.synchronize()
prof = h.profile_read()
print("%s H%d L%d B%d T%d %s: %.3f ms/step; per step ms: %s" % (model, H, L, B, T, prec, e0.elapsed_time(e1) / iters,
      {k: round(v[0] / iters, 3) for k, v in prof.items() if v[1]}), "loss", losses.cpu().numpy()[0])
