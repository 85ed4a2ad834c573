"""The reference's Python surface on the GPU: Model(args, infer) / Session.run call patterns of train.py and
sample.py, the state carry, checkpoints in save/<dataset>/, and the two scripts end to end on a synthetic
corpus (the kanji blob is not distributed with the reference)."""
import argparse
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import sketch_oracle as so

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _args(**kw):
    d = dict(rnn_size=32, num_layers=2, model="lstm", batch_size=6, seq_length=8, num_epochs=1, save_every=250,
             grad_clip=5.0, learning_rate=0.005, decay_rate=0.99, num_mixture=4, data_scale=15.0, keep_prob=1.0,
             stroke_importance_factor=200.0, dataset_name="toy")
    d.update(kw)
    return argparse.Namespace(**d)


def test_session_run_patterns_match_oracle():
    from sketch_rnn_b200.model import Model, Session
    args = _args()
    model = Model(args, precision="f32")
    sess = Session(model)
    cfg = so.Config(model="lstm", rnn_size=32, num_layers=2, num_mixture=4, batch_size=6, seq_length=8, keep_prob=1.0)
    params = {k: v.astype(np.float64) for k, v in model.handle.get_params_dict().items()}
    opt = so.AdamState(cfg, dtype=np.float64)
    x, y = so.synthetic_batch(6, 8, seed=3, p_eoc=0.1)
    sess.run(model.assign_lr(0.005 * 0.99 ** 2))
    state = sess.run(model.initial_state)
    assert len(state) == 2 and state[0].shape == (6, 64) and not state[0].any()
    st_ref = so.zero_state(cfg, 6, np.float64)
    for _ in range(2):  # state carried across batches (train.py:85-86)
        feed = {model.input_data: x, model.target_data: y, model.initial_state: state}
        loss, ls, lp, state, _ = sess.run([model.cost, model.cost_shape, model.cost_pen, model.final_state,
                                          model.train_op], feed)
        ref, st_ref, _ = so.train_step(cfg, params, opt, x.astype(np.float64), y.astype(np.float64), st_ref,
                                       0.005 * 0.99 ** 2)
        assert abs(loss - ref[0]) < 1e-4 * abs(ref[0]) and abs(lp - ref[2]) < 1e-4 * abs(ref[2])
        assert np.allclose(state[1], st_ref[1], atol=1e-5)
    out = sess.run([model.pi, model.sigma1, model.pen, model.final_state], {model.input_data: x, model.initial_state: state})
    fw = so.forward(cfg, params, x.astype(np.float64), [s.astype(np.float64) for s in state])
    assert np.allclose(out[0], fw["mdn"]["pi"], atol=1e-4) and np.allclose(out[1], fw["mdn"]["sigma1"], rtol=1e-4)


def test_infer_mutates_args_and_sample_signature():
    from sketch_rnn_b200.model import Model
    args = _args(batch_size=100, seq_length=300)
    model = Model(args, True)
    assert args.batch_size == 1 and args.seq_length == 1  # model.py:9-11
    np.random.seed(0)
    strokes, params = model.sample(None, 30, 0.5, 0.5, stop_if_eoc=False)
    assert strokes.shape == (30, 5) and strokes.dtype == np.float32 and len(params) == 30
    assert len(params[0]) == 7 and params[0][0].shape == (4,) and params[0][6].shape == (3,)
    assert np.all(strokes[:, 2:].sum(1) == 1)
    with pytest.raises(Exception, match="model type not supported: foo"):
        Model(_args(model="foo"))


def test_state_fed_back_in_place_equals_state_fed_back_as_copies():
    """train.py feeds the final_state of a step into the next one (train.py:84-86).  Model.train_step keeps that state in
    one buffer it updates in place; feeding copies (any other tuple) must give the same losses and states."""
    from sketch_rnn_b200.model import Model
    x, y = so.synthetic_batch(6, 8, seed=3)
    ma, mb = Model(_args(), precision="f32", seed=5), Model(_args(), precision="f32", seed=5)
    sa = sb = None
    for i in range(4):
        ca = ma.train_step(x, y, sa)
        cb = mb.train_step(x, y, sb)
        sa = ca[3]                                   # the returned tuple itself: in place
        sb = tuple(np.array(s) for s in cb[3])       # copies: stacked and copied in
        assert ca[:3] == cb[:3], (i, ca[:3], cb[:3])
        for u, v in zip(sa, sb):
            assert np.array_equal(u, v)
    # a fresh zero state after a run of steps is honoured (train.py:80 resets the state every epoch)
    ca = ma.train_step(x, y, ma.zero_state(6))
    cb = mb.train_step(x, y, None)
    assert ca[:3] == cb[:3]


def test_checkpoint_roundtrip(tmp_path):
    from sketch_rnn_b200.model import Model, Saver, Session, get_checkpoint_state
    args = _args()
    m1 = Model(args, precision="f32", seed=5)
    x, y = so.synthetic_batch(6, 8, seed=3)
    m1.lr = 0.004
    m1.train_step(x, y)
    d = tmp_path / "save" / "toy"
    d.mkdir(parents=True)
    Saver(m1).save(Session(m1), str(d / "model.ckpt"), global_step=7)
    ck = get_checkpoint_state(str(d))
    assert ck.model_checkpoint_path.endswith("model.ckpt-7")
    m2 = Model(_args(), precision="f32", seed=9)
    Saver(m2).restore(Session(m2), ck.model_checkpoint_path)
    v1, v2 = m1.variables(), m2.variables()
    assert set(v1) == set(v2) and "rnn_mdn/output_w/Adam_1" in v1 and "beta2_power" in v1
    for k in v1:
        assert np.array_equal(np.asarray(v1[k]), np.asarray(v2[k])), k
    a = m1.train_step(x, y)
    b = m2.train_step(x, y)
    assert a[0] == b[0]


def test_train_and_sample_scripts_end_to_end(tmp_path):
    from sketch_rnn_b200 import utils
    utils.make_synthetic_dataset(str(tmp_path / "data" / "toy.cpkl"), num_sketches=40, seed=1)
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "--dataset_name", "toy", "--rnn_size", "32",
                        "--num_mixture", "4", "--batch_size", "8", "--seq_length", "20", "--num_epochs", "2",
                        "--max_batches", "3"], cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "(epoch 0 batch 1), cost = " in r.stdout and "model saved to" in r.stdout
    assert os.path.exists(str(tmp_path / "save" / "toy" / "config.pkl"))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "sample.py"), "--dataset_name", "toy", "--num_picture", "2",
                        "--sample_length", "40", "--max_attempts", "30", "--filename", "out", "--temperature", "0.5"],
                       cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "loading model: " in r.stdout


def test_batched_rejection_sampling_accepts_what_the_one_by_one_loop_accepts(tmp_path):
    """sample.py draws candidates 64 per launch and filters the batch; the reference (sample.py:81-100) draws and
    tests one sketch at a time.  Under the same np.random state both keep the same sketches, bit for bit."""
    import sample as sample_script
    from sketch_rnn_b200.model import Model
    model = Model(_args(rnn_size=64, num_mixture=6, batch_size=100, seq_length=300), True, seed=3)
    p = model.handle.get_params_dict()
    p["rnn_mdn/output_b"][0:3] += np.array([1.5, -0.5, 0.0], dtype=np.float32)   # pen lifts often, a few ends of content
    model.handle.set_params_dict(p)
    num, temp = 60, 0.8
    # the reference's loop, one model.sample call per candidate, recording every candidate's statistics
    np.random.seed(11)
    singles = [model.sample(None, num, temp, temp, stop_if_eoc=True)[0] for _ in range(96)]
    n_eos = sorted(int(s[:, 2].sum()) for s in singles)
    lo, hi = n_eos[len(n_eos) // 3], n_eos[2 * len(n_eos) // 3] + 1      # rejects roughly a third on each side
    ok = [sample_script.accept_sketch(s, 0.0, 1e9, lo, hi) for s in singles]
    assert 0 < sum(ok) < len(ok), (lo, hi, n_eos)
    want = [s for s, a in zip(singles, ok) if a][:5]
    assert len(want) == 5
    for per_launch in (1, 7, 64):
        sargs = sample_script.build_parser().parse_args(
            ["--num_picture", "5", "--sample_length", str(num), "--temperature", str(temp), "--filename",
             str(tmp_path / ("out%d" % per_launch)), "--candidates_per_launch", str(per_launch), "--picture_size", "1e9"])
        np.random.seed(11)
        got, params = sample_script.sample_sketches(model, None, sargs, max_size_ratio=1.0, min_num_stroke=lo,
                                                    max_num_stroke=hi)
        assert len(got) == 5 and len(params) == 5 and len(params[0]) == got[0].shape[0]
        for g, w in zip(got, want):
            assert g.shape == w.shape and np.array_equal(g.view(np.uint32), w.view(np.uint32)), per_launch
        assert os.path.exists(str(tmp_path / ("out%d.svg" % per_launch)))


def test_saver_keeps_the_last_five_checkpoints(tmp_path):
    """tf.train.Saver's default max_to_keep = 5 (train.py:64,74 relies on it: a 500-epoch run saves hundreds of times)."""
    from sketch_rnn_b200.model import Model, Saver, Session, get_checkpoint_state
    m = Model(_args(), precision="f32")
    d = tmp_path / "save" / "toy"
    d.mkdir(parents=True)
    saver = Saver(m)
    for step in range(1, 8):
        saver.save(Session(m), str(d / "model.ckpt"), global_step=step * 250)
    left = sorted(f for f in os.listdir(str(d)) if f.startswith("model.ckpt-"))
    assert left == sorted("model.ckpt-%d" % (s * 250) for s in range(3, 8)), left
    idx = open(str(d / "checkpoint")).read()
    assert idx.count("all_model_checkpoint_paths") == 5 and 'model_checkpoint_path: "model.ckpt-1750"' in idx
    assert get_checkpoint_state(str(d)).model_checkpoint_path.endswith("model.ckpt-1750")
