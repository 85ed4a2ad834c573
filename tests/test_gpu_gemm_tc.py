"""The bf16 tcgen05/TMA GEMM building block (csrc/gemm_tc.cu) against torch.matmul on the same
bf16 operands, for every operand-major combination the training step uses, ragged sizes, bias,
and the split-K atomic epilogue."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _handle():
    from sketch_rnn_b200 import _lib
    return _lib.Handle(model="lstm", rnn_size=16, num_layers=2, num_mixture=3)


def _mk(rows, cols, seed):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(rows, cols, generator=g) * 0.5).to(torch.bfloat16).cuda()


CASES = [  # M, N, K
    (128, 128, 64), (128, 128, 256), (256, 384, 192), (300, 147, 256), (1000, 1024, 264), (77, 50, 40),
    # >= 148 output tiles with a 16-byte-aligned C: the persistent TMA-store variant (ragged M, N, K; 1 and many k-blocks)
    (5000, 520, 200), (19000, 128, 8), (2500, 1100, 520),
]


@pytest.mark.parametrize("a_k,b_k", [(True, True), (True, False), (False, True), (False, False)])
@pytest.mark.parametrize("M,N,K", CASES)
def test_gemm_matches_torch(a_k, b_k, M, N, K):
    import torch
    h = _handle()
    pad = lambda v: (v + 7) // 8 * 8
    # storage: K-major A is [M, K]; MN-major A is [K, M]  (row pitch padded to 8 elements for TMA)
    A_store = _mk(M, pad(K), 1)[:, :K] if a_k else _mk(K, pad(M), 1)[:, :M]
    B_store = _mk(N, pad(K), 2)[:, :K] if b_k else _mk(K, pad(N), 2)[:, :N]
    A = A_store.float() if a_k else A_store.float().t()
    B = B_store.float().t() if b_k else B_store.float()
    bias = torch.randn(N, device="cuda")
    ref = A @ B + bias
    c = h.debug_gemm_bf16(A_store, a_k, B_store, b_k, M, N, K, bias=bias)
    torch.cuda.synchronize()
    err = (c - ref).abs().max().item() / (ref.abs().max().item() + 1e-9)
    assert err < 2e-3, err
    h.close()


def test_split_k_accumulate():
    """Tall-skinny weight gradient shape: C[256, 1024] += X^T[256, R] * dY[R, 1024], R = 6000."""
    import torch
    h = _handle()
    R, M, N = 6000, 256, 1024
    X = _mk(R, M, 3)
    dY = _mk(R, N, 4)
    c0 = torch.randn(M, N, device="cuda")
    ref = c0 + X.float().t() @ dY.float()
    c = c0.clone()
    h.debug_gemm_bf16(X, False, dY, False, M, N, R, accumulate=True, c=c)
    torch.cuda.synchronize()
    err = (c - ref).abs().max().item() / ref.abs().max().item()
    assert err < 2e-3, err
    h.close()
