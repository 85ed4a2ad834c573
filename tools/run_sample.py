# one throughput-sampler call at BASELINE config 5's shape (for ncu): python tools/run_sample.py [n [steps]]
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sketch_rnn_b200 import _lib, utils
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
os.environ.setdefault("SKRNN_SAMPLER_GRAPH", "0")   # plain launches so that a profiler sees the kernels
h = _lib.Handle(model="lstm", rnn_size=256, num_layers=2, num_mixture=24, precision="bf16")
h.set_params(utils.default_init(h, seed=0))
for i in range(2):
    strokes, lengths, _ = h.sample(n, steps, 0.1, 0.1, False, seed=i)
torch.cuda.synchronize()
print("lengths", int(lengths.sum()))
