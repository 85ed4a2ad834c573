# per-step latency of the recurrent kernels vs batch size (profile API); usage: python tools/sweep_rnn.py H L
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sketch_rnn_b200 import _lib, utils
H = int(sys.argv[1]) if len(sys.argv) > 1 else 256
L = int(sys.argv[2]) if len(sys.argv) > 2 else 2
prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
T = 300
for B in [int(v) for v in (sys.argv[4].split(",") if len(sys.argv) > 4 else "16,64,100,128,256,512,800".split(","))]:
    h = _lib.Handle(model="lstm", rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=0.8, precision=prec)
    h.set_params(utils.default_init(h, seed=0))
    x, y = utils.synthetic_batch(B, T, seed=1)
    x = torch.from_numpy(x).cuda(); y = torch.from_numpy(y).cuda()
    st = torch.zeros((L, B, h.S), device="cuda")
    for i in range(2): h.loss_grad(x, y, st, seed=i)
    h.profile_enable(True)
    for i in range(3): h.loss_grad(x, y, st, seed=i)
    pr = h.profile_read()
    f, b_, g = pr["rnn_fwd"][0] / 3, pr["rnn_bwd"][0] / 3, (pr["gemm_tc"][0] + pr["gemm_f32"][0]) / 3
    print("H %d L %d B %4d: fwd %.3f ms (%.2f us/step/layer) bwd %.3f ms (%.2f us/step/layer) gemm %.3f ms" % (H, L, B, f, f * 1e3 / (T * L), b_, b_ * 1e3 / (T * L), g), flush=True)
    h.close()
