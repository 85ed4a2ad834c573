# throughput sampler (bf16 handle) timing: python tools/tp_time.py [n [steps]]; SKRNN_SAMPLE_TP=0 selects the step-by-step path
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sketch_rnn_b200 import _lib, utils
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
h = _lib.Handle(model="lstm", rnn_size=256, num_layers=2, num_mixture=24, precision="bf16")
h.set_params(utils.default_init(h, seed=0))
for i in range(2):
    strokes, lengths, _ = h.sample(n, steps, 0.1, 0.1, False, seed=i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(5):
    strokes, lengths, _ = h.sample(n, steps, 0.1, 0.1, False, seed=i)
e1.record()
torch.cuda.synchronize()
print("n=%d steps=%d TP=%s: %.3f ms per call (%d points)" % (n, steps, os.environ.get("SKRNN_SAMPLE_TP", "1"),
                                                            e0.elapsed_time(e1) / 5, int(lengths.sum())))
