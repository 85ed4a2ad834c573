"""skrnn_clip_adam_noalloc -- the optimiser step for callers that cannot allocate.

Some users of this library keep parameters, gradients and Adam slots in tensors that a framework
pool handed out, and the pool has usually claimed the whole GPU: a cudaMalloc from inside a library
call would fail or fragment it, so creating a handle (which allocates) is off the table.  The
no-alloc entry point works on their pointers and a fixed scratch array inside the library.  This
test drives it through the C ABI and checks it against the handle-based skrnn_apply_clip_adam on
the same numbers, for two consecutive calls (the second one proves that the shared scratch, which
still holds the first call's contents, is reset correctly), and checks that the size limit is
enforced and that device memory does not shrink across the call.  A second group checks it against
the CPU oracle's clip_by_global_norm + adam_apply (reference model.py:182-184) at the sizes that
take the kernel's other paths: config 4's 5,334,675 parameters (more chunks than blocks, a ragged
last chunk), exactly SKRNN_NOALLOC_MAX_PARAMS, and a small call right after a large one (stale
partial sums of the large call are still in the scratch array).

The file name sorts last on purpose: a run that stops at the first failure has then already
executed every other test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-4


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


def test_noalloc_matches_handle_entry_point_two_calls_in_a_row():
    import torch
    from sketch_rnn_b200 import _lib
    h = _lib.Handle(model="lstm", rnn_size=64, num_layers=2, num_mixture=8)
    n = h.P
    g = torch.Generator(device="cpu").manual_seed(0)
    p0 = torch.randn(n, generator=g) * 0.1
    h.set_params(p0.numpy())
    p = p0.clone().cuda()
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    gn2 = torch.zeros(1, device="cuda")
    b1, b2 = 0.9, 0.999
    # step 1: big gradient (clips), step 2: small gradient (does not clip)
    for step, scale in enumerate([3.0, 1e-3]):
        grad = torch.randn(n, generator=g) * scale
        h.grads_tensor().copy_(grad.cuda())
        gn1 = h.apply_clip_adam(0.005)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        gd = grad.cuda()
        torch.cuda.synchronize()
        free_mid = torch.cuda.mem_get_info()[0]
        _lib.clip_adam_noalloc(gd, p, m, v, 0.005, 5.0, b1, b2, gnorm_out=gn2)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_mid <= free_before
        b1 *= 0.9
        b2 *= 0.999
        assert abs(float(gn1) - float(gn2)) <= TOL * float(gn1), (step, float(gn1), float(gn2))
        ref_p = torch.from_numpy(h.get_params()).cuda()
        ref_m, ref_v, hb1, hb2 = h.get_adam_state()
        assert _rel(p, ref_p) < TOL, step
        assert _rel(m, torch.from_numpy(ref_m).cuda()) < TOL, step
        assert _rel(v, torch.from_numpy(ref_v).cuda()) < TOL, step
        assert abs(hb1 - b1) < 1e-12 and abs(hb2 - b2) < 1e-12
    h.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from sketch_rnn_b200 import _lib
    t = torch.zeros(8, device="cuda")
    lib = _lib.load()
    rc = lib.skrnn_clip_adam_noalloc(_lib.NOALLOC_MAX_PARAMS + 1, t.data_ptr(), t.data_ptr(), t.data_ptr(),
                                     t.data_ptr(), 0.005, 5.0, 0.9, 0.999, None, None)
    assert rc != 0
    assert b"exceed the fixed scratch capacity" in lib.skrnn_last_error()
    rc = lib.skrnn_clip_adam_noalloc(0, t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 0.005, 5.0,
                                     0.9, 0.999, None, None)
    assert rc != 0


class _Opt(object):
    """The oracle's Adam state for one flat tensor (oracle.sketch_oracle.AdamState without a model config)."""

    def __init__(self, n):
        self.m = {"p": np.zeros(n, dtype=np.float32)}
        self.v = {"p": np.zeros(n, dtype=np.float32)}
        self.beta1_power = 0.9
        self.beta2_power = 0.999


def _oracle_vs_noalloc(sizes_and_scales, seed):
    """Runs the calls in order on the GPU (fresh tensors per size) and the oracle beside them; two steps per size so
    the slots and beta powers of the second step are non-trivial."""
    import torch
    from oracle import sketch_oracle as so
    from sketch_rnn_b200 import _lib
    rng = np.random.RandomState(seed)
    for n, scales in sizes_and_scales:
        p_ref = {"p": (rng.standard_normal(n) * 0.1).astype(np.float32)}
        opt = _Opt(n)
        p = torch.from_numpy(p_ref["p"].copy()).cuda()
        m = torch.zeros(n, device="cuda")
        v = torch.zeros(n, device="cuda")
        gn_dev = torch.zeros(1, device="cuda")
        for scale in scales:
            g = (rng.standard_normal(n) * scale).astype(np.float32)
            clipped, gn_ref = so.clip_by_global_norm({"p": g}, 5.0)
            b1, b2 = opt.beta1_power, opt.beta2_power
            so.adam_apply(p_ref, clipped, opt, 0.005)
            _lib.clip_adam_noalloc(torch.from_numpy(g).cuda(), p, m, v, 0.005, 5.0, b1, b2, gnorm_out=gn_dev)
            torch.cuda.synchronize()
            assert abs(float(gn_dev) - float(gn_ref)) <= TOL * float(gn_ref), (n, scale, float(gn_dev), float(gn_ref))
            for got, ref, name in ((p, p_ref["p"], "p"), (m, opt.m["p"], "m"), (v, opt.v["p"], "v")):
                e = _rel(got.cpu(), torch.from_numpy(ref))
                assert e < TOL, (n, scale, name, e)
        del p, m, v
        torch.cuda.empty_cache()


def test_noalloc_against_oracle_at_config4_size_then_a_small_call():
    # 5,334,675 = config 4's parameter count: 1303 chunks on ~592 blocks (grid-stride loop), last chunk ragged;
    # then 1000 parameters: one chunk, while partial[1..1302] still hold the large call's sums
    _oracle_vs_noalloc([(5334675, [3.0, 1e-4]), (1000, [2.0, 1e-3]), (4097, [1.0])], seed=1)


def test_noalloc_against_oracle_at_the_size_limit():
    from sketch_rnn_b200 import _lib
    _oracle_vs_noalloc([(_lib.NOALLOC_MAX_PARAMS, [0.5])], seed=2)
