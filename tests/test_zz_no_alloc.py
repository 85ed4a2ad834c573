"""skrnn_clip_adam_noalloc -- the optimiser step for callers that cannot allocate.

Some users of this library keep parameters, gradients and Adam slots in tensors that a framework
pool handed out, and the pool has usually claimed the whole GPU: a cudaMalloc from inside a library
call would fail or fragment it, so creating a handle (which allocates) is off the table.  The
no-alloc entry point works on their pointers and a fixed scratch array inside the library.  This
test drives it through the C ABI and checks it against the handle-based skrnn_apply_clip_adam on
the same numbers, for two consecutive calls (the second one proves that the shared scratch, which
still holds the first call's contents, is reset correctly), and checks that the size limit is
enforced and that device memory does not shrink across the call.

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
