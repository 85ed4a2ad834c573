"""CPU tests of the oracle itself (no GPU): manual BPTT vs torch fp64 autograd, finite differences,
golden vectors, the C sampler restatement vs the NumPy transcription, reference quirks."""
import glob
import os
import pickle

import numpy as np
import pytest

from oracle import sketch_oracle as so
from oracle import sketch_oracle_torch as st
from oracle import sampler_ref

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _case(model, H=8, L=2, M=3, B=3, T=6, seed=0, p_eoc=0.2):
    cfg = so.Config(model=model, rnn_size=H, num_layers=L, num_mixture=M, batch_size=B, seq_length=T, keep_prob=0.8)
    p = so.init_params(cfg, seed=seed + 1, dtype=np.float64, trained_like=True)
    x, y = so.synthetic_batch(B, T, seed=seed + 2, dtype=np.float64, p_eoc=p_eoc)
    s0 = so.synthetic_state(cfg, B, dtype=np.float64)
    mask = so.synthetic_dropout_mask(cfg, B, T).astype(np.float64)
    return cfg, p, x, y, s0, mask


@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
@pytest.mark.parametrize("L", [2, 3])
def test_manual_backward_matches_torch_autograd(model, L):
    cfg, p, x, y, s0, mask = _case(model, L=L)
    l1, g1, f1, m1 = so.loss_and_grads(cfg, p, x, y, s0, mask)
    l2, g2, f2, m2 = st.loss_and_grads(cfg, p, x, y, s0, mask)
    assert np.allclose(l1, l2, rtol=1e-12)
    for n in g1:
        assert np.allclose(g1[n], g2[n], rtol=1e-9, atol=1e-12), n
    for a, b in zip(f1, f2):
        assert np.allclose(a, b, atol=1e-13)
    for k in ["pi", "mu1", "sigma1", "corr", "pen"]:
        assert np.allclose(m1[k], m2[k], atol=1e-13)


@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
def test_finite_differences(model):
    cfg, p, x, y, s0, mask = _case(model, H=8, B=2, T=5)
    _, g, _, _ = so.loss_and_grads(cfg, p, x, y, s0, mask)
    rng = np.random.RandomState(0)
    for name in p:
        for _ in range(3):
            idx = tuple(rng.randint(0, s) for s in p[name].shape)
            old = p[name][idx]
            eps = 1e-6
            p[name][idx] = old + eps
            lp = so.forward(cfg, p, x, s0, y, mask)["losses"][0]
            p[name][idx] = old - eps
            lm = so.forward(cfg, p, x, s0, y, mask)["losses"][0]
            p[name][idx] = old
            fd = (lp - lm) / (2 * eps)
            assert abs(fd - g[name][idx]) < 1e-5 * max(1.0, abs(fd)), (name, idx, fd, g[name][idx])


@pytest.mark.parametrize("path", GOLDEN)
def test_golden_vectors(path):
    z = np.load(path)
    cfg = so.Config(model=str(z["model"]), rnn_size=int(z["H"]), num_layers=int(z["L"]), num_mixture=int(z["M"]),
                    batch_size=int(z["B"]), seq_length=int(z["T"]), keep_prob=0.8)
    params = so.unflatten(cfg, z["params"].astype(np.float64))
    s0 = [s.astype(np.float64) for s in z["state_in"]]
    losses, grads, fstate, _ = so.loss_and_grads(cfg, params, z["x"].astype(np.float64), z["y"].astype(np.float64),
                                                 s0, z["mask"].astype(np.float64))
    assert np.allclose(losses, z["losses"], rtol=1e-10)
    assert np.allclose(so.flatten(cfg, grads, np.float64), z["grads"], rtol=1e-8, atol=1e-12)
    assert np.allclose(np.stack(fstate, 0), z["final_state"], atol=1e-12)
    # fp32 evaluation of the oracle stays within the fp32-path tolerance of its own fp64 answer
    p32 = so.unflatten(cfg, z["params"])
    l32, g32, _, _ = so.loss_and_grads(cfg, p32, z["x"], z["y"], [s for s in z["state_in"]], z["mask"])
    assert abs(l32[0] - z["losses"][0]) < 1e-4 * abs(z["losses"][0])
    gs = np.max(np.abs(z["grads"]))
    assert np.max(np.abs(so.flatten(cfg, g32, np.float64) - z["grads"])) < 1e-4 * gs
    # sampler trace
    strokes, mixp = sampler_ref.sample(cfg, p32, z["sample_uniforms"], z["sample_uniforms"].shape[0], 0.4, 0.4,
                                       False, want_params=True)
    assert np.array_equal(strokes.view(np.uint32), z["sample_strokes"].view(np.uint32))
    assert np.array_equal(mixp.view(np.uint32), z["sample_mixp"].view(np.uint32))


def test_eoc_reset_restores_fed_state_not_zeros():
    """model.py:82-92: the reset source is the FED initial_state; the resetting step's output is
    computed before the reset."""
    cfg, p, x, y, s0, _ = _case("lstm", B=2, T=4, p_eoc=0.0)
    x[:, :, 2:5] = [0, 0, 1]
    x[0, 1, 2:5] = [0, 1, 0]  # row 0: eoc on the input of step 1
    outs, fstate, _ = so.unroll_fwd(cfg, p, x, s0)
    # state after step 1 for row 0 equals the fed state, so steps 2.. replay from s0
    x2 = x.copy()
    outs2, _, _ = so.unroll_fwd(cfg, p, x2[:, 2:], s0)
    assert np.allclose(outs[0, 2:], outs2[0])
    assert not np.allclose(outs[1, 2:], outs2[1])
    # an all-eoc sequence leaves final_state == fed state
    x[:, :, 2:5] = [0, 1, 0]
    _, fs, _ = so.unroll_fwd(cfg, p, x, s0)
    for a, b in zip(fs, s0):
        assert np.array_equal(a, b)


def test_clamped_rows_have_zero_shape_gradient():
    """-log(max(S, 1e-20)): a row whose mixture density underflows contributes log(1e20) and no
    gradient through the mixture (model.py:130)."""
    cfg, p, x, y, s0, _ = _case("lstm", B=2, T=3)
    y[:, :, 0:2] = 1e4  # absurdly far target -> density underflows to 0
    fw = so.forward(cfg, p, x, s0, y)
    assert np.isclose(fw["losses"][1], -np.log(1e-20))
    g = {n: np.zeros_like(v) for n, v in p.items()}
    _, dl = so.head_bwd(cfg, p, g, fw["_hcache"], 2, 3)
    assert np.all(dl[:, 3:] == 0) and np.any(dl[:, :3] != 0)


def test_clip_and_adam_formulas():
    cfg = so.Config(model="rnn", rnn_size=8, num_layers=2, num_mixture=2)
    p = so.init_params(cfg, seed=0, dtype=np.float64)
    g = {n: np.ones_like(v) for n, v in p.items()}
    P = so.param_count(cfg)
    clipped, gn = so.clip_by_global_norm(g, 5.0)
    assert np.isclose(gn, np.sqrt(P))
    assert np.isclose(so.global_norm(clipped), 5.0)
    small = {n: v * 1e-3 for n, v in g.items()}
    c2, gn2 = so.clip_by_global_norm(small, 5.0)
    assert all(np.allclose(c2[n], small[n]) for n in small)  # gn < clip: untouched
    opt = so.AdamState(cfg, dtype=np.float64)
    before = {n: v.copy() for n, v in p.items()}
    so.adam_apply(p, clipped, opt, 0.01)
    # first step: m = 0.1 g, v = 0.001 g^2, alpha = lr*sqrt(1-b2)/(1-b1); epsilon 1e-3 un-corrected
    gg = clipped["rnn_mdn/output_b"][0]
    alpha = 0.01 * np.sqrt(1 - 0.999) / (1 - 0.9)
    want = before["rnn_mdn/output_b"][0] - alpha * 0.1 * gg / (np.sqrt(0.001 * gg * gg) + 1e-3)
    assert np.isclose(p["rnn_mdn/output_b"][0], want)
    assert np.isclose(opt.beta1_power, 0.81) and np.isclose(opt.beta2_power, 0.999 ** 2)


@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
def test_c_sampler_matches_numpy_transcription(model):
    cfg = so.Config(model=model, rnn_size=16, num_layers=2, num_mixture=5)
    p = so.init_params(cfg, seed=3, trained_like=True)
    u = np.random.RandomState(1).rand(40, 4).astype(np.float32)
    for tm, stop in [(1.0, False), (0.3, False), (0.1, True)]:
        s1, m1 = so.sample(cfg, p, u, 40, tm, tm, stop)
        s2, m2 = sampler_ref.sample(cfg, p, u, 40, tm, tm, stop, want_params=True)
        assert s1.shape == s2.shape
        assert np.array_equal(s1[:, 2:], s2[:, 2:])
        assert np.max(np.abs(s1[:, :2] - s2[:, :2])) < 1e-4 * cfg.data_scale
        ref_pi = np.stack([m[0] for m in m1])
        assert np.max(np.abs(ref_pi - m2[:, 3:3 + 5])) < 1e-4 * max(1.0, 1.0 / tm)


def test_sampler_quirks():
    """temperature only from the third point, pen temperature never applied, pi_pdf/temp_pen in the
    returned params, get_pi_idx == -1 -> last component / cont, eoc input holds the state."""
    cfg = so.Config(model="lstm", rnn_size=16, num_layers=2, num_mixture=4)
    p = so.init_params(cfg, seed=5, trained_like=True)
    u = np.random.RandomState(2).rand(6, 4).astype(np.float32)
    _, mp_hot = sampler_ref.sample(cfg, p, u, 6, 1.0, 1.0, False, want_params=True)
    _, mp_cold = sampler_ref.sample(cfg, p, u, 6, 0.25, 0.25, False, want_params=True)
    assert np.array_equal(mp_hot[:2, :3 + 4], mp_cold[:2, :3 + 4])      # i = 0, 1: no temperature
    assert np.allclose(mp_cold[2:, 3:7].sum(1), 1.0 / 0.25, rtol=1e-5)  # i > 1: pi_pdf / temp_pen
    assert np.allclose(mp_cold[:, :3].sum(1), 1.0, rtol=1e-6)           # pen pdf is a plain softmax
    u_hi = u.copy()
    u_hi[:, 0:2] = 1.5  # cumsum never reaches u -> idx -1
    s, _ = sampler_ref.sample(cfg, p, u_hi, 6, 1.0, 1.0, False, want_params=True)
    assert np.all(s[:, 4] == 1) and np.all(s[:, 2:4] == 0)
    s_np, _ = so.sample(cfg, p, u_hi, 6, 1.0, 1.0, False)
    assert np.array_equal(s_np[:, 2:], s[:, 2:])


def test_shipped_config_pickle_loads():
    """save/kanji/config.pkl (py2 protocol-0 pickle of argparse.Namespace, train.py:54-55) -- a copy
    of its 514 bytes is not vendored; the loader is exercised on an equivalent pickle we write."""
    import argparse
    ns = argparse.Namespace(grad_clip=5.0, rnn_size=256, data_scale=15.0, learning_rate=0.005, num_layers=2,
                            seq_length=300, decay_rate=0.99, num_mixture=24, batch_size=100, num_epochs=500,
                            dataset_name="kanji", model="lstm", save_every=250, keep_prob=0.8,
                            stroke_importance_factor=200.0)
    blob = pickle.dumps(ns, protocol=0)
    back = pickle.loads(blob, encoding="latin1")
    cfg = so.Config(model=back.model, rnn_size=back.rnn_size, num_layers=back.num_layers,
                    num_mixture=back.num_mixture)
    assert so.param_count(cfg) == 831379
    assert so.param_count(so.Config(model="lstm", rnn_size=512, num_layers=3)) == 5334675
    assert so.param_count(so.Config(model="gru")) == 632979
    assert so.param_count(so.Config(model="rnn")) == 236179


def test_unknown_model_raises_reference_message():
    with pytest.raises(Exception, match="model type not supported: foo"):
        so.Config(model="foo")


def test_shared_math_header_against_libm():
    """oracle/sampler_ref.c and the CUDA sampler share csrc/skrnn_math.h (own exp / log / tanh / sincos so that host
    and device produce the same bits).  Bit-equality between the two therefore says nothing about whether skm_exp IS
    exp: this test measures the header's functions against libm (NumPy float64) over the ranges the sampler feeds
    them -- softmax / sigma exponents, log of probabilities, tanh of rho-hat and cell states, Box-Muller angles."""
    import ctypes as C
    lib = sampler_ref.lib()
    lib.skref_sincos2pi.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    rng = np.random.RandomState(0)

    def worst_rel(fn, ref, xs):
        got = np.array([fn(float(v)) for v in xs], dtype=np.float64)
        want = ref(xs.astype(np.float64))
        return float(np.max(np.abs(got - want) / np.abs(want)))

    xs = np.concatenate([rng.uniform(-87, 88, 20000), rng.uniform(-1, 1, 5000)]).astype(np.float32)
    assert worst_rel(lib.skref_exp, np.exp, xs) < 2e-7          # ~1.3 ulp of float32
    xs = np.concatenate([np.exp(rng.uniform(-87, 88, 20000)), rng.uniform(0.5, 2, 5000)]).astype(np.float32)
    xs = xs[np.abs(xs - 1.0) > 1e-3]                             # relative error of log is ill-defined at log(1) = 0
    assert worst_rel(lib.skref_log, np.log, xs) < 3e-7
    xs = np.concatenate([rng.uniform(-10, 10, 20000), rng.uniform(-0.01, 0.01, 5000)]).astype(np.float32)
    got = np.array([lib.skref_tanh(float(v)) for v in xs], dtype=np.float64)
    assert np.max(np.abs(got - np.tanh(xs.astype(np.float64)))) < 1.5e-7
    s, c = C.c_float(), C.c_float()
    worst = 0.0
    for u in rng.uniform(0, 1, 5000).astype(np.float32):
        lib.skref_sincos2pi(float(u), C.byref(s), C.byref(c))
        worst = max(worst, abs(s.value - np.sin(2 * np.pi * float(u))), abs(c.value - np.cos(2 * np.pi * float(u))))
    assert worst < 2e-7
    # saturation instead of 0 / inf at the ends of the float32 range (documented behaviour of the header)
    assert 0.0 < lib.skref_exp(-200.0) < 2e-38 and 1e38 < lib.skref_exp(100.0) < np.inf
    assert lib.skref_tanh(50.0) == 1.0 and lib.skref_tanh(-50.0) == -1.0
