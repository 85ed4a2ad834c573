"""GPU parity: the CUDA path (through the C ABI in include/skrnn.h) against the CPU oracle on the
same seeded inputs.  Tolerances are the ones BASELINE.json's north_star states: loss / gradients
1e-4 relative on the fp32 path, MDN parameters 1e-4, sampled strokes bit-exact."""
import numpy as np
import pytest

from oracle import sketch_oracle as so
from oracle import sampler_ref

pytestmark = pytest.mark.gpu

TOL = 1e-4


def rel_err(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(a - ref)) / (np.max(np.abs(ref)) + 1e-30))


def make_handle(cfg, params, B=0, T=0, precision="f32"):
    from sketch_rnn_b200 import _lib
    h = _lib.Handle(model=cfg.model, rnn_size=cfg.rnn_size, num_layers=cfg.num_layers,
                    num_mixture=cfg.num_mixture, batch_size=B, seq_length=T, keep_prob=cfg.keep_prob,
                    grad_clip=cfg.grad_clip, stroke_importance_factor=cfg.stroke_importance_factor,
                    data_scale=cfg.data_scale, precision=precision)
    h.set_params_dict(params)
    return h


def stack_state(state):
    return np.ascontiguousarray(np.stack(state, 0), dtype=np.float32)


def f64(params):
    return {k: v.astype(np.float64) for k, v in params.items()}


CASES = [
    # model, H, L, M, B, T
    ("lstm", 16, 2, 3, 5, 7),
    ("lstm", 32, 3, 4, 19, 9),
    ("gru", 16, 2, 3, 5, 7),
    ("gru", 64, 3, 5, 11, 6),
    ("rnn", 16, 2, 3, 5, 7),
    ("rnn", 40, 3, 24, 9, 8),
    ("lstm", 256, 2, 24, 23, 12),
    ("gru", 256, 2, 24, 10, 10),
    ("lstm", 512, 2, 24, 9, 5),
]


def _case(model, H, L, M, B, T, seed=0, keep=0.8):
    cfg = so.Config(model=model, rnn_size=H, num_layers=L, num_mixture=M, batch_size=B, seq_length=T,
                    keep_prob=keep)
    params = so.init_params(cfg, seed=seed + 1, trained_like=True)
    x, y = so.synthetic_batch(B, T, seed=seed + 2, p_eoc=0.15)
    x[0, 0, 2:5] = [0, 1, 0]   # eoc on the very first step of row 0
    x[-1, -1, 2:5] = [0, 1, 0]  # and on the very last step of the last row
    s0 = so.synthetic_state(cfg, B, seed=seed)
    mask = so.synthetic_dropout_mask(cfg, B, T, seed=seed)
    return cfg, params, x, y, s0, mask


@pytest.mark.parametrize("model,H,L,M,B,T", CASES)
def test_forward_mdn_and_state(model, H, L, M, B, T):
    cfg, params, x, y, s0, _ = _case(model, H, L, M, B, T)
    ref = so.forward(cfg, f64(params), x.astype(np.float64), [s.astype(np.float64) for s in s0])
    h = make_handle(cfg, params, B, T)
    mdn, state_out = h.forward(x, stack_state(s0))
    mdn = mdn.cpu().numpy()
    state_out = state_out.cpu().numpy()
    Mx = cfg.num_mixture
    names = ["pi", "mu1", "mu2", "sigma1", "sigma2", "corr"]
    assert rel_err(mdn[:, 0:3], ref["mdn"]["pen"]) < TOL
    for i, n in enumerate(names):
        got = mdn[:, 3 + i * Mx:3 + (i + 1) * Mx]
        assert rel_err(got, ref["mdn"][n]) < TOL, n
    for l in range(L):
        assert rel_err(state_out[l], ref["final_state"][l]) < TOL
    h.close()


@pytest.mark.parametrize("model,H,L,M,B,T", CASES)
def test_loss_and_gradients(model, H, L, M, B, T):
    cfg, params, x, y, s0, mask = _case(model, H, L, M, B, T)
    losses_ref, grads_ref, fstate_ref, _ = so.loss_and_grads(
        cfg, f64(params), x.astype(np.float64), y.astype(np.float64),
        [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    h = make_handle(cfg, params, B, T)
    losses, state_out = h.loss_grad(x, y, stack_state(s0), dropout_mask=mask)
    losses = losses.cpu().numpy()
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    for i in range(3):
        assert abs(losses[i] - losses_ref[i]) <= TOL * abs(losses_ref[i]) + 1e-6, (i, losses, losses_ref)
    errs = {n: rel_err(g[n], grads_ref[n]) for n in grads_ref}  # per tensor, against that tensor's own scale
    assert max(errs.values()) < TOL, errs
    so_state = state_out.cpu().numpy()
    for l in range(L):
        assert rel_err(so_state[l], fstate_ref[l]) < TOL
    h.close()


def test_sharded_loss_matches_full_batch():
    """Two data-parallel shards (rows 0..B/2, B/2..B) summed == the full batch."""
    cfg, params, x, y, s0, mask = _case("lstm", 32, 2, 4, 12, 6)
    h = make_handle(cfg, params, 12, 6)
    st = stack_state(s0)
    losses, _ = h.loss_grad(x, y, st, dropout_mask=mask)
    g_full = h.grads_tensor().cpu().numpy().copy()
    l_full = losses.cpu().numpy()
    acc_g, acc_l = 0, 0
    for r in range(2):
        sl = slice(6 * r, 6 * r + 6)
        lo, _ = h.loss_grad(x[sl], y[sl], np.ascontiguousarray(st[:, sl]), dropout_mask=mask[sl], row0=6 * r,
                            loss_rows=12 * 6)
        acc_g = acc_g + h.grads_tensor().cpu().numpy()
        acc_l = acc_l + lo.cpu().numpy()
    assert rel_err(acc_g, g_full) < TOL
    assert rel_err(acc_l, l_full) < TOL
    h.close()


def test_in_kernel_dropout_is_shard_invariant():
    """In-kernel dropout is keyed by the global row, so sharding does not change the mask."""
    cfg, params, x, y, s0, _ = _case("lstm", 32, 2, 4, 8, 6)
    h = make_handle(cfg, params, 8, 6)
    st = stack_state(s0)
    lo, _ = h.loss_grad(x, y, st, seed=1234)
    g_full = h.grads_tensor().cpu().numpy().copy()
    acc = 0
    for r in range(2):
        sl = slice(4 * r, 4 * r + 4)
        h.loss_grad(x[sl], y[sl], np.ascontiguousarray(st[:, sl]), seed=1234, row0=4 * r, loss_rows=48)
        acc = acc + h.grads_tensor().cpu().numpy()
    assert rel_err(acc, g_full) < TOL
    lo2, _ = h.loss_grad(x, y, st, seed=99)
    assert abs(float(lo2[0]) - float(lo[0])) > 0  # a different seed draws a different mask
    h.close()


@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
def test_train_steps_clip_adam(model):
    """Three sess.run(train_op) steps with state carry (train.py:85-86): weights, Adam slots and
    losses track the oracle; the first step clips (gn > 5), later ones may not."""
    cfg, params, x, y, s0, mask = _case(model, 32, 2, 4, 6, 8)
    p64 = f64(params)
    opt = so.AdamState(cfg, dtype=np.float64)
    h = make_handle(cfg, params, 6, 8)
    st_ref = [s.astype(np.float64) for s in s0]
    st = stack_state(s0)
    import torch
    for step in range(3):
        lr = 0.005 * (0.99 ** step)
        losses_ref, st_ref, gn_ref = so.train_step(cfg, p64, opt, x.astype(np.float64), y.astype(np.float64),
                                                   st_ref, lr, mask.astype(np.float64))
        losses, st_t = h.loss_grad(x, y, st, dropout_mask=mask)
        gn = h.apply_clip_adam(lr)
        torch.cuda.synchronize()
        st = st_t.cpu().numpy()
        assert abs(float(gn) - gn_ref) <= TOL * gn_ref
        assert abs(float(losses[0]) - losses_ref[0]) <= 5 * TOL * abs(losses_ref[0])
    got = h.get_params_dict()
    for n in p64:
        assert rel_err(got[n], p64[n]) < 5 * TOL, n
    m, v, b1, b2 = h.get_adam_state()
    assert abs(b1 - opt.beta1_power) < 1e-12 and abs(b2 - opt.beta2_power) < 1e-12
    md = h.get_params_dict(m)
    for n in p64:
        assert rel_err(md[n], opt.m[n]) < 5 * TOL, n
    h.close()


def test_train_step_host_entry_point():
    cfg, params, x, y, s0, _ = _case("lstm", 32, 2, 4, 6, 8, keep=1.0)
    p64 = f64(params)
    opt = so.AdamState(cfg, dtype=np.float64)
    losses_ref, st_ref, _ = so.train_step(cfg, p64, opt, x.astype(np.float64), y.astype(np.float64),
                                          [s.astype(np.float64) for s in s0], 0.005)
    h = make_handle(cfg, params, 6, 8)
    st = stack_state(s0)
    losses = h.train_step_host(x, y, st, 0.005)
    assert rel_err(losses, losses_ref) < TOL
    for l in range(2):
        assert rel_err(st[l], st_ref[l]) < TOL
    got = h.get_params_dict()
    for n in p64:
        assert rel_err(got[n], p64[n]) < 5 * TOL, n
    h.close()


@pytest.mark.parametrize("kernel", ["cluster", "streaming"])
@pytest.mark.parametrize("model,H,L,M", [("lstm", 16, 2, 3), ("gru", 32, 2, 5), ("rnn", 24, 3, 4),
                                         ("lstm", 256, 2, 24), ("gru", 256, 2, 24), ("rnn", 64, 3, 7)])
def test_sampler_bit_exact(model, H, L, M, kernel, monkeypatch):
    """Sampled stroke sequences are bit-identical to the CPU restatement when both are fed the same
    uniforms; covers temperature, stop_if_eoc and several sketches per CTA -- through the cluster-resident
    kernel (sampler_cl.cu; rnn_size 24 is not a multiple of 16 and stays on the streaming kernel) and through
    the streaming kernel (sampler.cu)."""
    monkeypatch.setenv("SKRNN_SAMPLE_CLUSTER", "1" if kernel == "cluster" else "0")
    cfg = so.Config(model=model, rnn_size=H, num_layers=L, num_mixture=M)
    params = so.init_params(cfg, seed=7, trained_like=True)
    h = make_handle(cfg, params)
    steps = 40
    for n, temp, stop in [(1, 1.0, False), (1, 0.1, True), (7, 0.5, False), (21, 0.3, True)]:
        u = np.random.RandomState(n).rand(n, steps, 4).astype(np.float32)
        strokes, lengths, mixp = h.sample(n, steps, temp, temp, stop, uniforms=u, want_params=True)
        strokes, lengths, mixp = strokes.cpu().numpy(), lengths.cpu().numpy(), mixp.cpu().numpy()
        for i in range(n):
            ref_s, ref_p = sampler_ref.sample(cfg, params, u[i], steps, temp, temp, stop, want_params=True)
            k = ref_s.shape[0]
            assert lengths[i] == k, (i, lengths[i], k)
            assert np.array_equal(strokes[i, :k].view(np.uint32), ref_s.view(np.uint32)), (model, n, i)
            assert np.all(strokes[i, k:] == 0)
            assert np.array_equal(mixp[i, :k].view(np.uint32), ref_p.view(np.uint32))
    h.close()


@pytest.mark.parametrize("model,H,n", [("lstm", 256, 629), ("gru", 64, 629), ("lstm", 72, 2405)])
def test_sampler_bit_exact_many_sketches(model, H, n):
    """More sketches than 4 per SM: the streaming kernel's 16-sketches-per-CTA instance, ragged last CTA; a spread
    of sketches replayed by the CPU restatement."""
    cfg = so.Config(model=model, rnn_size=H, num_layers=2, num_mixture=24)
    params = so.init_params(cfg, seed=11, trained_like=True)
    h = make_handle(cfg, params)
    steps, temp = 14, 0.6
    u = np.random.RandomState(5).rand(n, steps, 4).astype(np.float32)
    strokes, lengths, _ = h.sample(n, steps, temp, temp, True, uniforms=u)
    strokes, lengths = strokes.cpu().numpy(), lengths.cpu().numpy()
    for i in [0, 1, 15, 16, 17, 300, n - 17, n - 2, n - 1]:
        ref_s, _ = sampler_ref.sample(cfg, params, u[i], steps, temp, temp, True, want_params=True)
        k = ref_s.shape[0]
        assert lengths[i] == k, (i, lengths[i], k)
        assert np.array_equal(strokes[i, :k].view(np.uint32), ref_s.view(np.uint32)), (model, i)
        assert np.all(strokes[i, k:] == 0)
    h.close()


@pytest.mark.parametrize("sb", [13, 14, 15, 16])
@pytest.mark.parametrize("model,H", [("lstm", 256), ("gru", 96), ("rnn", 40)])
def test_sampler_bit_exact_sketches_per_cta(monkeypatch, model, H, sb):
    """Every sketches-per-CTA instance of the streaming kernel (the library picks the one that fills the CTA slots of
    the GPU best; SKRNN_SAMPLE_SB pins it), ragged last CTA, rnn_size that is not a multiple of the copy group."""
    monkeypatch.setenv("SKRNN_SAMPLE_SB", str(sb))
    cfg = so.Config(model=model, rnn_size=H, num_layers=2, num_mixture=24)
    params = so.init_params(cfg, seed=13, trained_like=True)
    h = make_handle(cfg, params)
    n, steps, temp = 611, 9, 0.7
    u = np.random.RandomState(6).rand(n, steps, 4).astype(np.float32)
    strokes, lengths, mixp = h.sample(n, steps, temp, temp, False, uniforms=u, want_params=True)
    strokes, lengths, mixp = strokes.cpu().numpy(), lengths.cpu().numpy(), mixp.cpu().numpy()
    for i in [0, sb - 1, sb, 2 * sb + 1, 305, n - sb, n - 1]:
        ref_s, ref_p = sampler_ref.sample(cfg, params, u[i], steps, temp, temp, False, want_params=True)
        assert lengths[i] == steps
        assert np.array_equal(strokes[i].view(np.uint32), ref_s.view(np.uint32)), (model, sb, i)
        assert np.array_equal(mixp[i].view(np.uint32), ref_p.view(np.uint32)), (model, sb, i)
    h.close()


def test_sampler_against_numpy_transcription():
    """Independent check of the sampler against the NumPy transcription of model.py:187-264."""
    cfg = so.Config(model="lstm", rnn_size=32, num_layers=2, num_mixture=6)
    params = so.init_params(cfg, seed=11, trained_like=True)
    h = make_handle(cfg, params)
    u = np.random.RandomState(5).rand(3, 30, 4).astype(np.float32)
    strokes, lengths, _ = h.sample(3, 30, 0.4, 0.4, False, uniforms=u)
    strokes = strokes.cpu().numpy()
    for i in range(3):
        ref, _ = so.sample(cfg, params, u[i], 30, 0.4, 0.4, False)
        assert np.array_equal(strokes[i, :, 2:], ref[:, 2:])
        assert np.max(np.abs(strokes[i, :, :2] - ref[:, :2])) < 1e-3 * cfg.data_scale
    h.close()


def test_sampler_philox_seeded_mode_runs_and_is_deterministic():
    cfg = so.Config(model="lstm", rnn_size=32, num_layers=2, num_mixture=6)
    params = so.init_params(cfg, seed=11, trained_like=True)
    h = make_handle(cfg, params)
    a, la, _ = h.sample(9, 20, 0.5, 0.5, False, seed=42)
    b, lb, _ = h.sample(9, 20, 0.5, 0.5, False, seed=42)
    c, _, _ = h.sample(9, 20, 0.5, 0.5, False, seed=43)
    assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())
    assert not np.array_equal(a.cpu().numpy(), c.cpu().numpy())
    # sharding invariance: sketches 4..8 drawn as their own call with sketch0=4
    d, _, _ = h.sample(5, 20, 0.5, 0.5, False, seed=42, sketch0=4)
    assert np.array_equal(a.cpu().numpy()[4:], d.cpu().numpy())
    h.close()


def _softmax(z):
    e = np.exp(z - z.max(axis=1, keepdims=True))
    return e / e.sum(axis=1, keepdims=True)


@pytest.mark.parametrize("tp,cl,M,n", [("1", "1", 24, 192), ("1", "2", 24, 192), ("1", "4", 30, 150), ("0", "1", 24, 192)])
def test_throughput_sampler_bf16_replayed_by_oracle(monkeypatch, tp, cl, M, n):
    """tp = "1": the persistent kernel (sampler_tp.cu; 30 mixtures put the mixture logits across two 32-column pieces of
    the accumulator, 150 sketches leave the last tile ragged) with one CTA per 64-sketch tile (cl = "1") or the tile's gate
    columns split over a cluster of 2 or 4 CTAs; "0": the step-by-step kernels of sampler_tc.cu.
    The tensor-core sampler (bf16 handle, many sketches) is not bit-exact, so it is
    checked in two halves: (1) its per-step mixture parameters against the fp64 oracle replaying the
    sampler's OWN trajectory step by step (state hold on eoc included) within the bf16 tolerance 2e-2;
    (2) its draws against a NumPy redo of model.py:219-261 from the dumped parameters and uniforms."""
    monkeypatch.setenv("SKRNN_SAMPLE_TP", tp)
    monkeypatch.setenv("SKRNN_TP_CLUSTER", cl)
    cfg = so.Config(model="lstm", rnn_size=256, num_layers=2, num_mixture=M)
    params = so.init_params(cfg, seed=7, trained_like=True)
    h = make_handle(cfg, params, precision="bf16")
    steps, temp = 24, 0.6
    u = np.random.RandomState(3).rand(n, steps, 4).astype(np.float32)
    strokes, lengths, mixp = h.sample(n, steps, temp, temp, False, uniforms=u, want_params=True)
    strokes, lengths, mixp = strokes.cpu().numpy(), lengths.cpu().numpy(), mixp.cpu().numpy()
    assert np.all(lengths == steps)
    assert np.all(strokes[:, :, 2:].sum(axis=2) == 1.0)
    assert strokes[:, :, 3].sum() > 0, "no eoc drawn: the state-hold branch is not covered"
    p64 = f64(params)
    state = so.zero_state(cfg, n, np.float64)
    x = np.zeros((n, 1, 5))
    worst = {}
    for i in range(steps):
        fw = so.forward(cfg, p64, x, state)
        mdn = fw["mdn"]
        got = mixp[:, i]
        pi = mdn["pi"]
        if i > 1:
            pi = _softmax(np.log(pi) / temp) / temp
        ref = {"pen": _softmax(mdn["pen"]), "pi": pi, "mu1": mdn["mu1"], "mu2": mdn["mu2"],
               "sigma1": mdn["sigma1"], "sigma2": mdn["sigma2"], "corr": mdn["corr"]}
        blocks = {"pen": got[:, 0:3]}
        for k, name in enumerate(["pi", "mu1", "mu2", "sigma1", "sigma2", "corr"]):
            blocks[name] = got[:, 3 + k * M:3 + (k + 1) * M]
        for name in ref:
            worst[name] = max(worst.get(name, 0.0), rel_err(blocks[name], ref[name]))
        # (2) redo the draw from the dumped parameters
        p = blocks["pi"] * (temp if i > 1 else 1.0)
        cdf = np.cumsum(p.astype(np.float32), axis=1)
        hit = cdf >= u[:, i, 0:1]
        idx = np.where(hit.any(axis=1), hit.argmax(axis=1), M - 1)
        pen = blocks["pen"]
        pen_idx = np.where(pen[:, 0] >= u[:, i, 1], 0, np.where(pen[:, 0] + pen[:, 1] >= u[:, i, 1], 1, 2))
        r = np.arange(n)
        z1, z2 = so.box_muller(u[:, i, 2], u[:, i, 3])
        rho = blocks["corr"][r, idx]
        x1 = blocks["mu1"][r, idx] + blocks["sigma1"][r, idx] * z1
        x2 = blocks["mu2"][r, idx] + blocks["sigma2"][r, idx] * (rho * z1 + np.sqrt(1 - rho * rho) * z2)
        ok_pen = strokes[r, i, 2 + pen_idx] == 1.0
        ok_xy = (np.abs(strokes[:, i, 0] - x1 * cfg.data_scale) < 1e-3 * cfg.data_scale) & \
                (np.abs(strokes[:, i, 1] - x2 * cfg.data_scale) < 1e-3 * cfg.data_scale)
        # draws on a CDF boundary may round either way; allow a handful
        assert ok_pen.mean() > 0.98 and ok_xy.mean() > 0.98, (i, ok_pen.mean(), ok_xy.mean())
        # the sampler's own point is the next input (model.py:259-261)
        x = np.zeros((n, 1, 5))
        x[:, 0, 0:2] = strokes[:, i, 0:2] / cfg.data_scale
        x[:, 0, 2:5] = strokes[:, i, 2:5]
        state = fw["final_state"]
    assert max(worst.values()) < 2e-2, worst
    h.close()


@pytest.mark.parametrize("H,L", [(128, 3), (256, 2)])
def test_throughput_sampler_stop_determinism_and_sharding(H, L):
    """(128, 3): the step-by-step kernels; (256, 2): the persistent kernel, whose CTAs own 64 sketches each (160 sketches:
    a ragged last CTA; the shard of 96 starts in the middle of what was a CTA of the full call)."""
    cfg = so.Config(model="lstm", rnn_size=H, num_layers=L, num_mixture=8)
    params = so.init_params(cfg, seed=5, trained_like=True)
    h = make_handle(cfg, params, precision="bf16")
    n, steps = 160, 30
    u = np.random.RandomState(9).rand(n, steps, 4).astype(np.float32)
    full, _, _ = h.sample(n, steps, 0.8, 0.8, False, uniforms=u)
    cut, lens, _ = h.sample(n, steps, 0.8, 0.8, True, uniforms=u)
    full, cut, lens = full.cpu().numpy(), cut.cpu().numpy(), lens.cpu().numpy()
    for i in range(n):
        eoc = np.nonzero(full[i, :, 3])[0]
        k = eoc[0] + 1 if len(eoc) else steps
        assert lens[i] == k
        assert np.array_equal(cut[i, :k], full[i, :k]) and np.all(cut[i, k:] == 0)
    assert (lens < steps).any()
    a, _, _ = h.sample(n, steps, 0.5, 0.5, False, seed=42)
    b, _, _ = h.sample(n, steps, 0.5, 0.5, False, seed=42)
    c, _, _ = h.sample(n, steps, 0.5, 0.5, False, seed=43)
    assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())
    assert not np.array_equal(a.cpu().numpy(), c.cpu().numpy())
    d, _, _ = h.sample(96, steps, 0.5, 0.5, False, seed=42, sketch0=64)
    assert np.array_equal(a.cpu().numpy()[64:], d.cpu().numpy())
    h.close()


@pytest.mark.parametrize("n,steps,stop,cl", [(64, 1, False, ""), (65, 2, False, ""), (127, 3, True, ""), (9600, 5, False, ""),
                                               (20000, 4, True, ""), (64, 301, False, ""), (3000, 7, True, ""),
                                               (700, 9, False, "1"), (700, 9, False, "2"), (9600, 3, False, "2")])
def test_persistent_sampler_edge_shapes(monkeypatch, n, steps, stop, cl):
    """sample_tp_kernel at the edges of its launch shape: one step, a one-row second tile, more CTAs than SMs (two and
    three waves), an odd number of steps beyond 300, the sketch counts that select clusters of 4 (up to 2368 sketches), 2
    (up to 4736) and single CTAs by themselves (cl = "") or forced; and the same sketches drawn again as the tail of a
    70-sketch call (another tile, other rows of it, possibly another cluster size) are bit-identical."""
    if cl:
        monkeypatch.setenv("SKRNN_TP_CLUSTER", cl)
    cfg = so.Config(model="lstm", rnn_size=256, num_layers=2, num_mixture=24)
    params = so.init_params(cfg, seed=21, trained_like=True)
    h = make_handle(cfg, params, precision="bf16")
    s, l, _ = h.sample(n, steps, 0.7, 0.7, stop, seed=3)
    s, l = s.cpu().numpy(), l.cpu().numpy()
    assert np.all(np.isfinite(s)) and np.all((l >= 1) & (l <= steps))
    if not stop:
        assert np.all(l == steps)
    for i in (0, n // 2, n - 1):
        k = l[i]
        assert np.all(s[i, :k, 2:].sum(axis=1) == 1.0) and np.all(s[i, k:] == 0)
        if stop and k < steps:
            assert s[i, k - 1, 3] == 1.0
    if n >= 70:
        s2, l2, _ = h.sample(70, steps, 0.7, 0.7, stop, seed=3, sketch0=n - 70)
        assert np.array_equal(s2.cpu().numpy(), s[n - 70:]) and np.array_equal(l2.cpu().numpy(), l[n - 70:])
    h.close()


def test_errors_are_reported_not_thrown():
    from sketch_rnn_b200 import _lib
    with pytest.raises(Exception, match="model type not supported"):
        _lib.Handle(model="transformer")
    with pytest.raises(_lib.SkrnnError, match="multiple of 8"):
        _lib.Handle(model="lstm", rnn_size=30)
    h = _lib.Handle(model="lstm", rnn_size=16, num_layers=2, num_mixture=3)
    with pytest.raises(_lib.SkrnnError, match="expected"):
        h.set_params(np.zeros(5, dtype=np.float32))
    h.close()


BF16_TOL = 2e-2
BF16_CASES = [("lstm", 16, 2, 3, 5, 7), ("gru", 64, 3, 5, 11, 6), ("rnn", 40, 3, 24, 9, 8),
              ("lstm", 256, 2, 24, 23, 12), ("gru", 256, 2, 24, 10, 10), ("lstm", 512, 3, 24, 9, 5),
              # gru / rnn tensor-core kernels: one CTA per batch tile (H <= 128), the cluster of two with an uneven
              # split of the units (H = 192: 128 + 64), several ragged 16-row tiles
              ("gru", 128, 2, 24, 37, 9), ("rnn", 128, 2, 24, 16, 7), ("gru", 192, 2, 24, 21, 8),
              ("rnn", 192, 3, 24, 33, 6), ("rnn", 256, 2, 24, 50, 9), ("gru", 256, 3, 24, 49, 7),
              # lstm, 32-row tiles on 16-CTA clusters: forward two halves of 16 rows with h pushed rank to rank, BPTT on
              # the single-tile kernel with pushed partial sums (the shapes a 4- or 8-GPU shard of config 4 runs)
              ("lstm", 512, 2, 24, 130, 10), ("lstm", 512, 2, 24, 70, 9)]


@pytest.mark.parametrize("model,H,L,M,B,T", BF16_CASES)
def test_bf16_path_loss_and_gradients(model, H, L, M, B, T):
    """bf16 tensor-core path (tcgen05 GEMMs): loss and gradients within 2e-2 of the fp64 oracle,
    MDN parameters within 2e-2 as well (its inputs went through bf16 products)."""
    cfg, params, x, y, s0, mask = _case(model, H, L, M, B, T)
    losses_ref, grads_ref, fstate_ref, _ = so.loss_and_grads(
        cfg, f64(params), x.astype(np.float64), y.astype(np.float64),
        [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    h = make_handle(cfg, params, B, T, precision="bf16")
    losses, state_out = h.loss_grad(x, y, stack_state(s0), dropout_mask=mask)
    losses = losses.cpu().numpy()
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    for i in range(3):
        assert abs(losses[i] - losses_ref[i]) <= BF16_TOL * abs(losses_ref[i]) + 1e-5, (i, losses, losses_ref)
    errs = {n: rel_err(g[n], grads_ref[n]) for n in grads_ref}  # per tensor, against that tensor's own scale
    assert max(errs.values()) < BF16_TOL, errs
    ref = so.forward(cfg, f64(params), x.astype(np.float64), [s.astype(np.float64) for s in s0])
    mdn, _ = h.forward(x, stack_state(s0))
    mdn = mdn.cpu().numpy()
    merr = {"pen": rel_err(mdn[:, 0:3], ref["mdn"]["pen"])}
    for i, n in enumerate(["pi", "mu1", "mu2", "sigma1", "sigma2", "corr"]):
        merr[n] = rel_err(mdn[:, 3 + i * M:3 + (i + 1) * M], ref["mdn"][n])
    assert max(merr.values()) < BF16_TOL, merr
    h.close()


@pytest.mark.parametrize("H,L,B,T", [(32, 1, 1, 1), (64, 2, 3, 2), (128, 2, 17, 5), (256, 1, 129, 3), (256, 2, 300, 4),
                                     (512, 1, 70, 3), (512, 2, 260, 2), (256, 1, 530, 3), (512, 1, 460, 3),
                                     # two-half-tile kernels: one timestep, a second half with a single row, every
                                     # cluster size that can multicast (2, 4, 8, 16 CTAs), exactly full tiles
                                     (512, 1, 460, 1), (512, 1, 33, 2), (256, 2, 64, 3), (128, 1, 40, 3), (64, 1, 40, 2),
                                     (512, 1, 128, 2)])
def test_bf16_tcgen05_recurrent_edge_shapes(H, L, B, T):
    """Shapes that exercise the tcgen05 recurrent kernels' corner cases: one timestep, one row, ragged last
    cluster, several clusters, every cluster size (1..16 CTAs), all batch tiles (16..128 columns; 64 and 128
    run the two-half-tile forward kernel)."""
    cfg, params, x, y, s0, mask = _case("lstm", H, L, 4, B, T, seed=3)
    x[:, :, 2:5] = [0, 0, 1]
    x[::3, T // 2, 2:5] = [0, 1, 0]  # eoc on a third of the rows in the middle of the sequence
    x[0, 0, 2:5] = [0, 1, 0]
    losses_ref, grads_ref, fstate_ref, _ = so.loss_and_grads(
        cfg, f64(params), x.astype(np.float64), y.astype(np.float64),
        [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    h = make_handle(cfg, params, B, T, precision="bf16")
    losses, state_out = h.loss_grad(x, y, stack_state(s0), dropout_mask=mask)
    losses = losses.cpu().numpy()
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    assert abs(losses[0] - losses_ref[0]) <= BF16_TOL * abs(losses_ref[0]) + 1e-5, (losses, losses_ref)
    errs = {n: rel_err(g[n], grads_ref[n]) for n in grads_ref}
    assert max(errs.values()) < BF16_TOL, errs
    so_state = state_out.cpu().numpy()
    for l in range(L):
        assert rel_err(so_state[l], fstate_ref[l]) < BF16_TOL
    h.close()


def test_bf16_in_kernel_dropout_is_shard_invariant():
    cfg, params, x, y, s0, _ = _case("lstm", 256, 2, 4, 40, 6)
    h = make_handle(cfg, params, 40, 6, precision="bf16")
    st = stack_state(s0)
    h.loss_grad(x, y, st, seed=77)
    g_full = h.grads_tensor().cpu().numpy().copy()
    acc = 0
    for r in range(2):
        sl = slice(20 * r, 20 * r + 20)
        h.loss_grad(x[sl], y[sl], np.ascontiguousarray(st[:, sl]), seed=77, row0=20 * r, loss_rows=240)
        acc = acc + h.grads_tensor().cpu().numpy()
    assert rel_err(acc, g_full) < 2e-3  # same mask, same bf16 roundings row by row; only sum order differs
    h.close()


def test_bf16_training_tracks_fp32_over_many_steps():
    """Twenty train steps (loss_grad + clip + Adam, state carried, in-kernel dropout) on the bf16 tensor-core path and
    on the fp32 path from the same initialisation: the loss trajectories and the final weights stay together.  This
    runs the whole step orchestration repeatedly (side-stream weight gradients, d(pre) double buffering, the
    interleaved weight copies being refreshed after every update)."""
    import torch
    cfg = so.Config(model="lstm", rnn_size=256, num_layers=3, num_mixture=8, batch_size=72, seq_length=24, keep_prob=0.9)
    params = so.init_params(cfg, seed=4, trained_like=True)
    x, y = so.synthetic_batch(72, 24, seed=5, p_eoc=0.05)
    traj = {}
    final = {}
    for prec in ("f32", "bf16"):
        h = make_handle(cfg, params, 72, 24, precision=prec)
        st = stack_state(so.zero_state(cfg, 72))
        ls = []
        for step in range(20):
            losses, st_t = h.loss_grad(x, y, st, seed=100 + step)
            h.apply_clip_adam(0.005)
            st = st_t
            ls.append(float(losses[0]))
        torch.cuda.synchronize()
        traj[prec] = np.array(ls)
        final[prec] = h.get_params_dict()
        h.close()
    assert traj["f32"][-1] < traj["f32"][0], "training did not reduce the loss"
    assert np.max(np.abs(traj["bf16"] - traj["f32"]) / np.abs(traj["f32"])) < 2e-2, (traj["bf16"], traj["f32"])
    for n in final["f32"]:
        d = np.max(np.abs(final["bf16"][n] - final["f32"][n])) / (np.max(np.abs(final["f32"][n])) + 1e-12)
        assert d < 5e-2, (n, d)


# ---------------------------------------------------------------------------------------------------------------
# committed known-answer vectors (tests/golden/*.npz) and the edge cases of model.py:130 / model.py:197 on the GPU
# ---------------------------------------------------------------------------------------------------------------
import glob
import os

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_golden_vectors_on_gpu(path):
    """The fixtures test_oracle.py pins the oracle with, fed to the CUDA path: losses, every gradient tensor, final
    state and MDN parameters within 1e-4 (fp32 path), the sampler trace bit for bit."""
    from sketch_rnn_b200 import _lib
    z = np.load(path)
    model, H, L, M, B, T = str(z["model"]), int(z["H"]), int(z["L"]), int(z["M"]), int(z["B"]), int(z["T"])
    h = _lib.Handle(model=model, rnn_size=H, num_layers=L, num_mixture=M, batch_size=B, seq_length=T, keep_prob=0.8)
    h.set_params(z["params"])
    losses, state_out = h.loss_grad(z["x"], z["y"], z["state_in"], dropout_mask=z["mask"])
    losses = losses.cpu().numpy()
    for i in range(3):
        assert abs(losses[i] - z["losses"][i]) <= TOL * abs(z["losses"][i]) + 1e-7
    g = h.grads_tensor().cpu().numpy()
    for name, off, shape in h.tensors():
        k = int(np.prod(shape))
        assert rel_err(g[off:off + k], z["grads"][off:off + k]) < TOL, name
    assert rel_err(state_out.cpu().numpy(), z["final_state"]) < TOL
    mdn, st = h.forward(z["x"], z["state_in"])
    mdn = mdn.cpu().numpy()
    assert rel_err(mdn[:, 0:3], z["mdn_nodrop"][:, 0:3]) < TOL
    for i in range(6):
        sl = slice(3 + i * M, 3 + (i + 1) * M)
        assert rel_err(mdn[:, sl], z["mdn_nodrop"][:, sl]) < TOL, i
    assert rel_err(st.cpu().numpy(), z["final_state_nodrop"]) < TOL
    u = z["sample_uniforms"]
    strokes, lengths, mixp = h.sample(1, u.shape[0], 0.4, 0.4, False, uniforms=u[None], want_params=True)
    assert np.array_equal(strokes.cpu().numpy()[0].view(np.uint32), z["sample_strokes"].view(np.uint32))
    assert np.array_equal(mixp.cpu().numpy()[0].view(np.uint32), z["sample_mixp"].view(np.uint32))
    h.close()


@pytest.mark.parametrize("precision", ["f32", "bf16"])
def test_clamped_rows_on_gpu(precision):
    """model.py:130: rows whose mixture density underflows contribute -log(1e-20) and no mixture gradient.  Half the
    rows get an absurd target, so both branches of the clamp run in the same launch of mdn_head_kernel."""
    cfg, params, x, y, s0, mask = _case("lstm", 32, 2, 4, 6, 5)
    y[::2, :, 0:2] = 1e4
    losses_ref, grads_ref, _, _ = so.loss_and_grads(cfg, f64(params), x.astype(np.float64), y.astype(np.float64),
                                                    [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    assert losses_ref[1] > 0.4 * -np.log(1e-20)  # the clamp really is active in the oracle
    tol = TOL if precision == "f32" else BF16_TOL
    h = make_handle(cfg, params, 6, 5, precision=precision)
    losses, _ = h.loss_grad(x, y, stack_state(s0), dropout_mask=mask)
    losses = losses.cpu().numpy()
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    for i in range(3):
        assert abs(losses[i] - losses_ref[i]) <= tol * abs(losses_ref[i]), (i, losses, losses_ref)
    for n in grads_ref:
        assert rel_err(g[n], grads_ref[n]) < tol, (n, rel_err(g[n], grads_ref[n]))
    # all rows clamped: the mixture columns of the output layer receive exactly no gradient
    y[:, :, 0:2] = 1e4
    h.loss_grad(x, y, stack_state(s0), dropout_mask=mask)
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    assert np.all(g["rnn_mdn/output_b"][3:] == 0) and np.any(g["rnn_mdn/output_b"][:3] != 0)
    h.close()


def test_sampler_pi_idx_minus_one_on_gpu():
    """model.py:189-197: get_pi_idx returns -1 when the running sum never reaches the draw -> Python's negative index
    picks the LAST mixture component, and the pen falls through to `cont`.  u >= 1 forces it on every point."""
    cfg = so.Config(model="lstm", rnn_size=16, num_layers=2, num_mixture=4)
    params = so.init_params(cfg, seed=5, trained_like=True)
    h = make_handle(cfg, params)
    u = np.random.RandomState(2).rand(3, 8, 4).astype(np.float32)
    u[:, :, 0:2] = 1.5
    u[1, 3:, 0] = 0.3   # mixed: some points draw normally
    strokes, lengths, mixp = h.sample(3, 8, 1.0, 1.0, False, uniforms=u, want_params=True)
    strokes, mixp = strokes.cpu().numpy(), mixp.cpu().numpy()
    assert np.all(strokes[:, :, 4] == 1) and np.all(strokes[:, :, 2:4] == 0)
    for i in range(3):
        ref_s, ref_p = sampler_ref.sample(cfg, params, u[i], 8, 1.0, 1.0, False, want_params=True)
        assert np.array_equal(strokes[i].view(np.uint32), ref_s.view(np.uint32))
        ref_np, _ = so.sample(cfg, params, u[i], 8, 1.0, 1.0, False)
        assert np.array_equal(ref_np[:, 2:], strokes[i, :, 2:])
    # the point drawn with idx = -1 is centred on the last component: replay sketch 0, step 0 by hand
    M = 4
    p = mixp[0, 0]
    z1, z2 = so.box_muller(u[0, 0, 2], u[0, 0, 3])
    x1 = p[3 + M + (M - 1)] + p[3 + 3 * M + (M - 1)] * z1
    assert abs(strokes[0, 0, 0] - x1 * cfg.data_scale) < 1e-4 * cfg.data_scale
    h.close()


@pytest.mark.parametrize("precision", ["f32", "bf16"])
def test_in_kernel_dropout_keep_rate_and_independence(precision):
    """Keep rate and independence of the in-kernel dropout hash (the stand-in for TF's random_uniform in
    model.py:29-30).  A one-layer RNN with zero matrices and bias atanh(0.5) outputs 0.5 for every unit, row and step,
    and with output_w = 0 the logits do not depend on the mask; then
        d cost / d output_w[u, j] = sum over rows of 0.5 * mask[row, u] / keep * dlogits[row, j],
    and its ratio to the same gradient at keep_prob = 1 is the dlogits-weighted kept fraction of unit u divided by
    keep: expectation 1, different per unit, different per seed."""
    from sketch_rnn_b200 import _lib
    H, B, T, keep = 64, 64, 40, 0.8
    cfg = so.Config(model="rnn", rnn_size=H, num_layers=1, num_mixture=2, batch_size=B, seq_length=T, keep_prob=keep)
    params = so.init_params(cfg, seed=3)
    # h_t = tanh(b) = const for every unit, row and step: zero matrices, bias atanh(0.5)
    params["rnn_mdn/MultiRNNCell/Cell0/BasicRNNCell/Linear/Matrix"][:] = 0
    params["rnn_mdn/MultiRNNCell/Cell0/BasicRNNCell/Linear/Bias"][:] = np.arctanh(0.5)
    x, y = so.synthetic_batch(B, T, seed=1)
    st = np.zeros((1, B, H), dtype=np.float32)
    params["rnn_mdn/output_w"][:] = 0
    h1 = _lib.Handle(model="rnn", rnn_size=H, num_layers=1, num_mixture=2, batch_size=B, seq_length=T, keep_prob=1.0,
                     precision=precision)
    h1.set_params_dict(params)
    h1.loss_grad(x, y, st)
    g_all = h1.get_params_dict(h1.grads_tensor().cpu().numpy())["rnn_mdn/output_w"].astype(np.float64)
    h1.close()
    h = make_handle(cfg, params, B, T, precision=precision)
    gs = []
    for seed in (11, 12):
        h.loss_grad(x, y, st, seed=seed)
        gs.append(h.get_params_dict(h.grads_tensor().cpu().numpy())["rnn_mdn/output_w"].astype(np.float64))
    h.close()
    j = int(np.argmax(np.abs(g_all).sum(axis=0)))         # a logit column with a solid gradient
    ratios = np.stack([g[:, j] / g_all[:, j] for g in gs])   # [2 seeds, H units], expectation 1
    assert abs(ratios.mean() - 1.0) < 0.01, ratios.mean()
    assert ratios.std() > 1e-3                             # units really have different masks
    assert np.abs(ratios[0] - ratios[1]).max() > 1e-3       # and a new seed draws new masks
    c = np.corrcoef(ratios[0], ratios[1])[0, 1]
    assert abs(c) < 0.5, c                                  # seeds are not correlated unit by unit


@pytest.mark.parametrize("model,precision,H,L", [("lstm", "bf16", 256, 3), ("gru", "f32", 32, 2), ("rnn", "bf16", 64, 2)])
def test_gradient_segments_and_their_events(model, precision, H, L):
    """skrnn_grad_segment / skrnn_stream_wait_grad (the hooks of the overlapped data-parallel all-reduce): the
    segments tile the flat gradient in the order the backward pass finishes them, and a stream that waits for a
    segment's event reads that segment's final values although it never synchronised with the compute stream."""
    import torch
    cfg, params, x, y, s0, _ = _case(model, H, L, 4, 40, 6)
    h = make_handle(cfg, params, 40, 6, precision=precision)
    segs = h.grad_segments()
    assert [s[0] for s in segs] == [L] + list(range(L - 1, -1, -1))
    cover = sorted((o, c) for _, o, c in segs)
    assert cover[0][0] == 0 and all(cover[i][0] + cover[i][1] == cover[i + 1][0] for i in range(L)) and \
        cover[-1][0] + cover[-1][1] == h.P
    names = {n: (o, int(np.prod(s))) for n, o, s in h.tensors()}
    o, c = [(o, c) for i, o, c in segs if i == L][0]
    assert o == names["rnn_mdn/output_w"][0] and o + c == names["rnn_mdn/output_b"][0] + names["rnn_mdn/output_b"][1]
    side = torch.cuda.Stream()
    st = stack_state(s0)
    for rep in range(3):
        h.loss_grad(x, y, st, seed=rep)
        g = h.grads_tensor()
        parts = []
        for idx, off, cnt in segs:
            h.stream_wait_grad(idx, side)
            with torch.cuda.stream(side):
                parts.append((off, g[off:off + cnt].clone()))
        side.synchronize()
        torch.cuda.synchronize()
        final = g.cpu().numpy()
        for off, part in parts:
            assert np.array_equal(part.cpu().numpy(), final[off:off + part.numel()]), (rep, off)
    h.close()


def test_bf16_lstm_time_ranges_match_single_launch(monkeypatch):
    """Large batches run every LSTM layer as several launches over consecutive time ranges (so that the hoisted GEMMs
    overlap the neighbouring ranges' recurrence).  The ranged step must reproduce the single-launch step: same
    forward state bit for bit, same losses, gradients equal up to the summation order of the weight-gradient GEMMs;
    and both sit within the bf16 tolerance of the oracle."""
    cfg, params, x, y, s0, mask = _case("lstm", 512, 2, 4, 260, 66, seed=5)
    x[::5, 33, 2:5] = [0, 1, 0]   # eoc right at a range boundary region
    losses_ref, grads_ref, fstate_ref, _ = so.loss_and_grads(
        cfg, f64(params), x.astype(np.float64), y.astype(np.float64),
        [s.astype(np.float64) for s in s0], mask.astype(np.float64))
    res = {}
    for ranges in ("1", "3"):
        monkeypatch.setenv("SKRNN_TC_RANGES", ranges)
        h = make_handle(cfg, params, 260, 66, precision="bf16")
        losses, state_out = h.loss_grad(x, y, stack_state(s0), dropout_mask=mask)
        res[ranges] = (losses.cpu().numpy(), state_out.cpu().numpy(), h.get_params_dict(h.grads_tensor().cpu().numpy()),
                       h.launch_count())
        h.close()
    assert res["3"][3] > res["1"][3], "the ranged run did not launch more kernels: ranges were not taken"
    assert np.array_equal(res["1"][1], res["3"][1])
    # (the fused head adds the per-warp loss sums with fp32 atomics, so the last bits of a loss depend on arrival order)
    assert np.allclose(res["1"][0], res["3"][0], rtol=1e-5)
    for n in grads_ref:
        assert rel_err(res["3"][2][n], res["1"][2][n]) < 1e-3, n
        assert rel_err(res["3"][2][n], grads_ref[n]) < BF16_TOL, n
    for i in range(3):
        assert abs(res["3"][0][i] - losses_ref[i]) <= BF16_TOL * abs(losses_ref[i])
    assert rel_err(res["3"][1], np.stack(fstate_ref, 0)) < BF16_TOL
