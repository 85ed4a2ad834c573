# time the tcgen05 GEMM at the shapes of the C4 training step: python tools/gemm_sweep.py
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sketch_rnn_b200 import _lib

h = _lib.Handle(model="lstm", rnn_size=16, num_layers=2, num_mixture=3)
R = 240000
SHAPES = [  # name, M, N, K, a_kmajor, b_kmajor, accumulate
    ("xw_l0   [R,8]x[8,2048]", R, 2048, 8, True, False, False),
    ("xw_l1   [R,512]x[512,2048]", R, 2048, 512, True, False, False),
    ("logits  [R,512]x[512,147]", R, 147, 512, True, False, False),
    ("dout    [R,152]x[152,512]", R, 512, 152, True, True, False),
    ("dgrad   [R,2048]x[2048,512]", R, 512, 2048, True, True, False),
    ("wgrad   [512,R]x[R,2048]", 512, 2048, R, False, False, True),
    ("wgrad_o [512,R]x[R,147]", 512, 147, R, False, False, True),
]
for name, M, N, K, ak, bk, acc in SHAPES:
    pad = lambda v: (v + 7) // 8 * 8
    A = torch.randn((M, pad(K)) if ak else (K, pad(M)), device="cuda").to(torch.bfloat16)
    B = torch.randn((N, pad(K)) if bk else (K, pad(N)), device="cuda").to(torch.bfloat16)
    A = A[:, :K] if ak else A[:, :M]
    B = B[:, :K] if bk else B[:, :N]
    c = torch.zeros((M, N), device="cuda")
    for _ in range(2):
        h.debug_gemm_bf16(A, ak, B, bk, M, N, K, accumulate=acc, c=c)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        h.debug_gemm_bf16(A, ak, B, bk, M, N, K, accumulate=acc, c=c)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    fl = 2.0 * M * N * K
    byts = M * K * 2 + K * N * 2 + M * N * 4
    print("%-30s %8.3f ms  %7.1f TFLOP/s  %7.1f GB/s (min traffic)" % (name, ms, fl / ms / 1e9, byts / ms / 1e6))
