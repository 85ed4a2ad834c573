# Repeated train steps at several batch sizes (every tcgen05 recurrent kernel variant), checking that the loss stays
# finite and goes down and that nothing hangs: python tools/stress.py [steps]
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sketch_rnn_b200 import _lib, utils

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
for (H, L, B, T) in [(512, 3, 800, 300), (512, 3, 400, 300), (512, 3, 200, 300), (512, 3, 100, 300), (256, 2, 100, 300),
                     (256, 2, 777, 120), (128, 2, 300, 64), (64, 2, 90, 50)]:
    h = _lib.Handle(model="lstm", rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=0.8,
                    precision="bf16")
    h.set_params(utils.default_init(h, seed=0))
    x, y = utils.synthetic_batch(B, T, seed=1)
    x = torch.from_numpy(x).cuda(); y = torch.from_numpy(y).cuda()
    st = torch.zeros((L, B, h.S), device="cuda")
    t0 = time.time()
    first = last = None
    for i in range(steps):
        losses, st = h.loss_grad(x, y, st, seed=i)
        h.apply_clip_adam(0.002)
        if i == 0 or i == steps - 1 or i % 50 == 0:
            v = float(losses[0])
            assert np.isfinite(v), (H, L, B, T, i, v)
            first = v if first is None else first
            last = v
    torch.cuda.synchronize()
    print("H%d L%d B%d T%d: %d steps in %.2f s, loss %.4f -> %.4f" % (H, L, B, T, steps, time.time() - t0, first, last), flush=True)
    assert last < first
    h.close()
print("stress ok")
