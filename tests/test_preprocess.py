"""svg -> stroke preprocessing (reference utils.py:124-217 and the path grammar of its vendored svg.path):
hand-written SVG files with hand-computed expectations."""
import os

import numpy as np
import pytest

from sketch_rnn_b200 import svgpath
from sketch_rnn_b200.utils import SketchLoader

SVG = """<?xml version="1.0"?>
<svg xmlns="http://www.w3.org/2000/svg" width="109" height="109">
 <g id="strokes">
  <path id="s1" d="M10,10 L20,10 l0,10"/>
  <path id="s2" d="M 30 30 c 0,10 10,20 30,30 s 5,-2 8,-4"/>
 </g>
 <text>1</text>
</svg>
"""


def bezier(p0, p1, p2, p3, t):
    z = (1 - t) ** 3 * p0 + 3 * t * (1 - t) ** 2 * p1 + 3 * t * t * (1 - t) * p2 + t ** 3 * p3
    return z.real, z.imag


def test_path_grammar():
    seg = svgpath.segments
    assert seg("M10,10 L20,10 l0,10 z") == [("L", 10 + 10j, 20 + 10j), ("L", 20 + 10j, 20 + 20j), ("L", 20 + 20j, 10 + 10j)]
    # first moveto absolute even when lower case, further pairs are (relative) linetos, H/V, numbers running together
    assert seg("m5 5 10 0h5v-5H0") == [("L", 5 + 5j, 15 + 5j), ("L", 15 + 5j, 20 + 5j), ("L", 20 + 5j, 20 + 0j), ("L", 20 + 0j, 0j)]
    assert seg("M0 0L1.5-2 .5.5") == [("L", 0j, 1.5 - 2j), ("L", 1.5 - 2j, 0.5 + 0.5j)]
    # smooth cubic: reflection of the previous second control point; without a cubic before it, the current point
    s = seg("M 30 30 c 0,10 10,20 30,30 s 10,5 20,0")
    assert s[1] == ("C", 60 + 60j, 80 + 70j, 70 + 65j, 80 + 60j)
    assert seg("M1 1 S 2 2 3 3")[0] == ("C", 1 + 1j, 1 + 1j, 2 + 2j, 3 + 3j)
    assert seg("M0 0 Q1 1 2 0 T4 0") == [("Q", 0j, 1 + 1j, 2 + 0j), ("Q", 2 + 0j, 3 - 1j, 4 + 0j)]
    assert seg("M0 0 A 5 5 0 0 1 10 0")[0][:1] == ("A",) and seg("M0 0 a 5 5 0 0 1 10 0")[0][2] == 10 + 0j
    with pytest.raises(ValueError):
        seg("M0 0 L1")
    with pytest.raises(ValueError):
        seg("10 10 L 1 1")


def test_svg_to_strokes_matches_hand_computation(tmp_path):
    f = tmp_path / "k.svg"
    f.write_text(SVG)
    lens = []
    got = SketchLoader.svg_to_strokes(str(f), lens)
    # path 1: points (10,10) start, (20,10), (20,20 eos).  path 2: start (30,30); first cubic has chord
    # |(30,30)| = 42.4 -> n = 3 points at t = 1/3, 2/3, 1; second cubic (60,60)->(68,56): chord 8.9 -> n = 1 point, eos
    c1 = [bezier(30 + 30j, 30 + 40j, 40 + 50j, 60 + 60j, t) for t in (1 / 3., 2 / 3., 1.0)]
    c2 = bezier(60 + 60j, 80 + 70j, 65 + 58j, 68 + 56j, 1.0)
    absolute = [(10, 10, 0), (20, 10, 0), (20, 20, 1), (30, 30, 0)] + [(x, y, 0) for x, y in c1] + [(c2[0], c2[1], 1)]
    want = np.zeros((len(absolute) - 1, 4), dtype=np.float32)
    for i in range(1, len(absolute)):
        want[i - 1, 0] = absolute[i][0] - absolute[i - 1][0]
        want[i - 1, 1] = absolute[i][1] - absolute[i - 1][1]
        want[i - 1, 2] = absolute[i][2]
    want[-1, 3] = 1
    assert got.dtype == np.float32 and got.shape == want.shape
    assert np.allclose(got, want, atol=1e-5)
    assert np.allclose(lens, [10.0, 10.0, np.hypot(30, 30), np.hypot(8, 4)])


def test_loader_builds_cpkl_from_raw_svgs(tmp_path):
    raw = tmp_path / "data" / "tiny"
    (raw / "sub").mkdir(parents=True)
    for name in ("a.svg", "sub/b.svg"):
        (raw / name).write_text(SVG)
    (raw / "notes.txt").write_text("ignored")
    np.random.seed(0)
    ld = SketchLoader(batch_size=2, seq_length=5, scale_factor=15.0, data_filename="tiny", data_dir=str(tmp_path / "data"))
    assert os.path.exists(str(tmp_path / "data" / "tiny.cpkl"))
    assert ld.num_samples == 2 and len(ld.length_data) == 8
    x, y = ld.next_batch()
    assert x.shape == (2, 5, 5) and np.array_equal(x[:, 1:], y[:, :-1])
    # the second construction loads the file it wrote (py2-compatible protocol 2)
    ld2 = SketchLoader(batch_size=2, seq_length=5, scale_factor=15.0, data_filename="tiny", data_dir=str(tmp_path / "data"))
    assert np.allclose(ld2.raw_data[0], SketchLoader.svg_to_strokes(str(raw / "a.svg")) / np.array([15, 15, 1, 1], dtype=np.float32))
