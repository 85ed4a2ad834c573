"""SketchLoader (reference utils.py:105-283): the per-sketch batcher here must reproduce the reference's
per-point walk exactly.  The reference cannot be imported (Python 2, svgwrite, IPython), so the test
carries a direct restatement of that walk as the checker."""
import os
import pickle

import numpy as np
import pytest

from sketch_rnn_b200 import utils


def _reference_walk(raw, index, pointer, n, rs):
    """The reference's next_seq loop (utils.py:231-256), point by point."""
    out = np.zeros((n, 5), dtype=np.float32)
    sx = rs.rand() * 0.6 + 0.7
    sy = rs.rand() * 0.6 + 0.7
    idx = 0
    finished = False

    def tick(p):
        p += 1
        if p >= len(raw):
            return 0, True
        return p, False

    data = raw[index[pointer]]
    for i in range(n):
        out[i, 0:4] = data[idx]
        out[i, 4] = 0 if (out[i, 2] > 0 or out[i, 3] > 0) else 1
        idx += 1
        if idx >= len(data) - 1:
            out[i, 2:5] = (0, 1, 0)
            idx = 0
            pointer, f = tick(pointer)
            finished |= f
            data = raw[index[pointer]]
    pointer, f = tick(pointer)
    finished |= f
    out[:, 0] *= sx
    out[:, 1] *= sy
    return out, pointer, finished


@pytest.fixture
def corpus(tmp_path):
    path = utils.make_synthetic_dataset(str(tmp_path / "data" / "toy.cpkl"), num_sketches=37, seed=3)
    return str(tmp_path / "data"), path


def test_file_format_is_the_reference_cpkl(corpus):
    _, path = corpus
    raw = pickle.load(open(path, "rb"), encoding="latin1")
    assert isinstance(raw, list) and raw[0].dtype == np.float32 and raw[0].shape[1] == 4
    assert all(r[-1, 3] == 1 for r in raw)


@pytest.mark.parametrize("T,B", [(20, 4), (300, 7), (7, 50)])
def test_batches_equal_the_reference_walk(corpus, T, B):
    data_dir, path = corpus
    np.random.seed(11)
    loader = utils.SketchLoader(B, T, 15.0, "toy", data_dir=data_dir)
    # replay with an identical RandomState
    rs = np.random.RandomState(11)
    raw = [np.array(d, dtype=np.float32) for d in pickle.load(open(path, "rb"), encoding="latin1")]
    for d in raw:
        d[:, 0:2] /= 15.0
    index = rs.permutation(list(range(len(raw))))
    assert np.array_equal(index, loader.index)
    pointer = 0
    for _ in range(3):
        x, y = loader.next_batch()
        rows = []
        for _b in range(B):
            row, pointer, _ = _reference_walk(raw, index, pointer, T + 1, rs)
            rows.append(row)
        ref = np.array(rows, dtype=np.float32)
        assert np.array_equal(x, ref[:, :-1]) and np.array_equal(y, ref[:, 1:])
        assert loader.pointer == pointer
        assert x.shape == (B, T, 5) and x.dtype == np.float32
        assert np.all(x[:, :, 2:5].sum(-1) == 1)


def test_epoch_flag_and_reshuffle(corpus):
    data_dir, _ = corpus
    np.random.seed(0)
    loader = utils.SketchLoader(10, 50, 15.0, "toy", data_dir=data_dir)
    n = 0
    while not loader.epoch_finished:
        loader.next_batch()
        n += 1
    assert n >= 1
    before = list(loader.index)
    loader.reset_index_pointer()
    assert loader.pointer == 0 and not loader.epoch_finished
    assert sorted(loader.index) == sorted(before)


def test_missing_corpus_is_an_error(tmp_path):
    with pytest.raises(IOError):
        utils.SketchLoader(2, 5, 1.0, "nope", data_dir=str(tmp_path))


def test_calculate_start_point_and_svg(tmp_path):
    d = np.array([[10, 0, 0, 0, 1], [0, 10, 1, 0, 0], [-5, -5, 0, 1, 0]], dtype=np.float32)
    ax, ay, sx, sy = utils.calculate_start_point(d, 1.0, 200)
    assert (sx, sy) == (10.0, 10.0) and (ax, ay) == (95.0, 95.0)
    out = str(tmp_path / "o.svg")
    text = utils.draw_stroke_color_array([d, d], svg_filename=out, block_size=160, maxcol=5, color_mode=False)
    assert os.path.exists(out) and text.count("<path") == 6 and 'stroke="#000"' in text
    assert text.count(" L ") == 4  # the point after a pen lift is a bare move
    text2 = utils.draw_stroke_color(np.concatenate([d, d]), svg_filename=None)
    assert text2.count("<path") == 6


def test_prefetching_loader_yields_the_same_batches(tmp_path):
    """The background-thread loader hands out exactly what SketchLoader.next_batch would, epoch ends included."""
    from sketch_rnn_b200 import utils
    path = utils.make_synthetic_dataset(str(tmp_path / "data" / "toy.cpkl"), num_sketches=23, seed=4)
    runs = []
    for wrap in (False, True):
        np.random.seed(7)
        ld = utils.SketchLoader(4, 30, 15.0, "toy", data_dir=str(tmp_path / "data"))
        if wrap:
            ld = utils.PrefetchingLoader(ld)
        seen = []
        for epoch in range(3):
            ld.reset_index_pointer()
            while not ld.epoch_finished:
                x, y = ld.next_batch()
                seen.append((x.copy(), y.copy(), ld.pointer, ld.epoch_finished))
        runs.append(seen)
        if wrap:
            ld.close()
    assert len(runs[0]) == len(runs[1]) > 3
    for a, b in zip(*runs):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2:] == b[2:]
