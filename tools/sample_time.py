# latency of the reference-precision sampler for few sketches: python tools/sample_time.py [n [steps]]
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sketch_rnn_b200 import _lib, utils
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
model = sys.argv[3] if len(sys.argv) > 3 else "lstm"
h = _lib.Handle(model=model, rnn_size=256, num_layers=2, num_mixture=24, precision="f32")
h.set_params(utils.default_init(h, seed=0))
for mode in ("1", "0"):
    os.environ["SKRNN_SAMPLE_CLUSTER"] = mode
    for i in range(2):
        strokes, lengths, _ = h.sample(n, steps, 0.1, 0.1, False, seed=i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(5):
        strokes, lengths, _ = h.sample(n, steps, 0.1, 0.1, False, seed=i)
    e1.record()
    torch.cuda.synchronize()
    print("%s n=%d steps=%d cluster=%s: %.3f ms per call (%d points)" % (model, n, steps, mode, e0.elapsed_time(e1) / 5, int(lengths.sum())))
