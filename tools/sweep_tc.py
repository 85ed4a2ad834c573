# bring-up helper: which (lbo, sbo, adv) make each major combination correct?
import sys, itertools, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sketch_rnn_b200 import _lib
h = _lib.Handle(model="lstm", rnn_size=16, num_layers=2, num_mixture=3)
M, N, K = 256, 256, 128
g = torch.Generator().manual_seed(0)
mk = lambda r, c: (torch.randn(r, c, generator=g) * 0.5).to(torch.bfloat16).cuda()
for a_k, b_k in [(True, True), (False, True), (True, False), (False, False)]:
    A_s = mk(M, K) if a_k else mk(K, M)
    B_s = mk(N, K) if b_k else mk(K, N)
    A = A_s.float() if a_k else A_s.float().t()
    B = B_s.float().t() if b_k else B_s.float()
    ref = A @ B
    def run(ov):
        c = h.debug_gemm_bf16(A_s, a_k, B_s, b_k, M, N, K, overrides=ov)
        torch.cuda.synchronize()
        return ((c - ref).abs().max() / ref.abs().max()).item()
    e = run(None)
    print("majors A_k=%s B_k=%s default err %.4g" % (a_k, b_k, e), flush=True)
    if e > 1e-2:
        cands_mn = [(8192, 1024, 2048), (1024, 8192, 2048), (8192, 1024, 1024), (1024, 8192, 1024), (128, 1024, 2048), (1024, 128, 2048), (8192, 1024, 32), (16, 1024, 2048), (2048, 1024, 4096), (1024, 2048, 4096), (8192, 2048, 4096)]
        cands_k = [(0, 1024, 32), (16, 1024, 32), (1024, 1024, 32), (1, 1024, 32), (0, 1024, 16), (0, 512, 32)]
        for ca in (cands_k if a_k else cands_mn):
            for cb in (cands_k if b_k else cands_mn):
                ov = [ca[0] or 0, ca[1], ca[2], cb[0] or 0, cb[1], cb[2]]
                e = run(ov)
                if e < 1e-2:
                    print("   OK with A", ca, "B", cb, "err %.3g" % e, flush=True)
print("done")
