"""GPU parity at the shapes bench.py measures (BASELINE.json configs 1, 3, 4; seq_length 300).

The CUDA path (through the C ABI) against the fp64 oracle on the same seeded inputs, with the tolerances
BASELINE.json's north_star states, applied PER TENSOR: every gradient tensor's worst element error is measured
against that tensor's own largest element (1e-4 on the fp32 path, 2e-2 on the bf16 tensor-core path), the
three losses and the final state likewise, and all six MDN parameter blocks plus the pen logits.

Config 4 (lstm H512 L3 B800) costs minutes on the CPU oracle, so its 800 rows are 16 copies of 50 distinct
rows: batch rows are independent (model.py:73-94 never mixes them) and the losses are means over rows, so the
800-row step has exactly the losses and gradients of the 50-row step the oracle computes -- while the GPU runs
the very launch configuration bench.py times (7 clusters x 16 CTAs, two 64-row half tiles in flight).
"""
import numpy as np
import pytest

from oracle import sketch_oracle as so

pytestmark = pytest.mark.gpu

TOL = {"f32": 1e-4, "bf16": 2e-2}
MDN_NAMES = ["pi", "mu1", "mu2", "sigma1", "sigma2", "corr"]

_cache = {}


def f64(params):
    return {k: v.astype(np.float64) for k, v in params.items()}


def rel_err(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(a - ref)) / (np.max(np.abs(ref)) + 1e-30))


def oracle_case(model, H, L, B, T, keep=0.8, seed=0):
    """Seeded inputs and the fp64 oracle's answers for them (computed once per session)."""
    key = (model, H, L, B, T, keep, seed)
    if key in _cache:
        return _cache[key]
    cfg = so.Config(model=model, rnn_size=H, num_layers=L, num_mixture=24, batch_size=B, seq_length=T, keep_prob=keep)
    params = so.init_params(cfg, seed=seed + 1, trained_like=True)
    x, y = so.synthetic_batch(B, T, seed=seed + 2)
    x[0, 0, 2:5] = [0, 1, 0]     # eoc on the first step of a row, on the last step of another
    x[-1, -1, 2:5] = [0, 1, 0]
    s0 = so.synthetic_state(cfg, B, seed=seed)
    mask = so.synthetic_dropout_mask(cfg, B, T, seed=seed)
    p64, x64, y64 = f64(params), x.astype(np.float64), y.astype(np.float64)
    s64 = [s.astype(np.float64) for s in s0]
    losses, grads, fstate, _ = so.loss_and_grads(cfg, p64, x64, y64, s64, mask.astype(np.float64))
    fw = so.forward(cfg, p64, x64, s64)
    out = dict(cfg=cfg, params=params, x=x, y=y, s0=np.ascontiguousarray(np.stack(s0, 0), dtype=np.float32), mask=mask,
               losses=np.array(losses), grads=grads, fstate=np.stack(fstate, 0), mdn=fw["mdn"],
               fstate_nodrop=np.stack(fw["final_state"], 0))
    _cache[key] = out
    return out


def make_handle(cfg, params, B, T, precision):
    from sketch_rnn_b200 import _lib
    h = _lib.Handle(model=cfg.model, rnn_size=cfg.rnn_size, num_layers=cfg.num_layers, num_mixture=cfg.num_mixture,
                    batch_size=B, seq_length=T, keep_prob=cfg.keep_prob, grad_clip=cfg.grad_clip,
                    stroke_importance_factor=cfg.stroke_importance_factor, data_scale=cfg.data_scale,
                    precision=precision)
    h.set_params_dict(params)
    return h


def check_step(h, c, precision, x, y, s0, mask, copies=1):
    """loss_grad + forward on the handle against the oracle record `c`; returns {name: error} and the failures."""
    tol = TOL[precision]
    errs, bad = {}, []

    def note(name, err, limit=tol):
        errs[name] = err
        if not err < limit:
            bad.append((name, err, limit))

    losses, state_out = h.loss_grad(x, y, s0, dropout_mask=mask)
    losses = losses.cpu().numpy().astype(np.float64)
    for i, n in enumerate(["cost", "cost_shape", "cost_pen"]):
        note("loss/" + n, abs(losses[i] - c["losses"][i]) / abs(c["losses"][i]))
    g = h.get_params_dict(h.grads_tensor().cpu().numpy())
    for n, ref in c["grads"].items():
        note("grad/" + n, rel_err(g[n], ref))
    st = state_out.cpu().numpy()
    B0 = c["fstate"].shape[1]
    for l in range(st.shape[0]):
        for k in range(copies):
            note("state/%d/copy%d" % (l, k), rel_err(st[l, k * B0:(k + 1) * B0], c["fstate"][l]))
    mdn, st2 = h.forward(x, s0)
    mdn = mdn.cpu().numpy()
    M = c["cfg"].num_mixture
    T = x.shape[1]
    rows = mdn[:B0 * T]
    note("mdn/pen", rel_err(rows[:, 0:3], c["mdn"]["pen"]))
    for i, n in enumerate(MDN_NAMES):
        note("mdn/" + n, rel_err(rows[:, 3 + i * M:3 + (i + 1) * M], c["mdn"][n]))
    if copies > 1:  # the last copy's rows, too (another cluster, the ragged end of the batch)
        last = mdn[(copies - 1) * B0 * T:]
        note("mdn/pi/last_copy", rel_err(last[:, 3:3 + M], c["mdn"]["pi"]))
    return errs, bad


@pytest.mark.parametrize("precision", ["f32", "bf16"])
@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
def test_config1_and_config3_full_step_T300(model, precision):
    """BASELINE configs 1 (lstm) and 3 (gru, rnn): H256 L2 M24 B100 T300, keep_prob 0.8 (injected mask)."""
    c = oracle_case(model, 256, 2, 100, 300)
    h = make_handle(c["cfg"], c["params"], 100, 300, precision)
    errs, bad = check_step(h, c, precision, c["x"], c["y"], c["s0"], c["mask"])
    h.close()
    print(model, precision, {k: "%.2e" % v for k, v in errs.items()})
    assert not bad, bad


@pytest.mark.parametrize("precision", ["f32", "bf16"])
def test_config4_row_slice_T300(precision):
    """BASELINE config 4's model (lstm H512 L3 M24 T300) on 50 rows, both precisions."""
    c = oracle_case("lstm", 512, 3, 50, 300)
    h = make_handle(c["cfg"], c["params"], 50, 300, precision)
    errs, bad = check_step(h, c, precision, c["x"], c["y"], c["s0"], c["mask"])
    h.close()
    print("config4 slice", precision, {k: "%.2e" % v for k, v in errs.items()})
    assert not bad, bad


def test_config4_full_batch_launch_T300():
    """The benched launch itself: B=800 = 16 copies of the 50 oracle rows (see the module docstring), bf16 path."""
    c = oracle_case("lstm", 512, 3, 50, 300)
    copies = 16
    x = np.ascontiguousarray(np.tile(c["x"], (copies, 1, 1)))
    y = np.ascontiguousarray(np.tile(c["y"], (copies, 1, 1)))
    s0 = np.ascontiguousarray(np.tile(c["s0"], (1, copies, 1)))
    mask = np.ascontiguousarray(np.tile(c["mask"], (copies, 1, 1)))
    h = make_handle(c["cfg"], c["params"], 800, 300, "bf16")
    errs, bad = check_step(h, c, "bf16", x, y, s0, mask, copies=copies)
    launches = h.launch_count()
    h.close()
    print("config4 B=800", {k: "%.2e" % v for k, v in errs.items() if "copy" not in k or k.endswith(("copy0", "copy15"))},
          "launches", launches)
    assert not bad, bad


def test_config4_full_batch_launch_T300_bptt_v2(monkeypatch):
    """The same launch with the alternative BPTT kernel instance (weights in tensor memory, accumulator slots shared by
    the half tiles, partial sums fetched into shared memory by bulk copies): SKRNN_TC_BWD_V2=1."""
    monkeypatch.setenv("SKRNN_TC_BWD_V2", "1")
    test_config4_full_batch_launch_T300()


def test_bench_loss_check_matches_oracle_slice():
    """bench.py prints `loss_check`: its synthetic batch's first rows through the oracle.  The same helper is
    exercised here on a small shape so a drift between bench inputs and oracle inputs cannot go unnoticed."""
    import bench
    w = dict(kind="train", model="lstm", H=64, L=2, M=24, B=24, T=40)
    r = bench.loss_check(w, "bf16", rows=8)
    assert r["rel_err"] < 2e-2, r
    r = bench.loss_check(w, "f32", rows=8)
    assert r["rel_err"] < 1e-4, r
