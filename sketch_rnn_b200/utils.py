"""Data path and drawing helpers with the reference's names and behaviour (reference utils.py).

`SketchLoader` keeps the reference semantics line for line (utils.py:105-283) -- same `.cpkl`
format, same batch layout, same order of `np.random` calls -- so a given NumPy seed produces the
same batches as the Python 2 original would.  The SVG writer emits the same document structure
as `draw_stroke_color_array` (utils.py:32-83) as plain text (svgwrite / IPython are not needed).
"""
import os
import pickle
import random

import numpy as np


def calculate_start_point(data, factor=1.0, block_size=200):
    """Reference utils.py:10-30: running-sum bounding box with per-step round(.., 3)."""
    sx = sy = 0
    maxx = minx = maxy = miny = 0
    for i in range(len(data)):
        sx += round(float(data[i, 0]) * factor, 3)
        sy += round(float(data[i, 1]) * factor, 3)
        maxx = max(maxx, sx)
        minx = min(minx, sx)
        maxy = max(maxy, sy)
        miny = min(miny, sy)
    abs_x = block_size / 2 - (maxx - minx) / 2 - minx
    abs_y = block_size / 2 - (maxy - miny) / 2 - miny
    return abs_x, abs_y, (maxx - minx), (maxy - miny)


def draw_stroke_color_array(data, factor=1, svg_filename="sample.svg", stroke_width=1, block_size=200,
                            maxcol=5, svg_only=False, color_mode=True):
    """Reference utils.py:32-83 as a plain-text SVG writer (svgwrite / IPython are not needed): a white
    rectangle, then for every sketch j (cell j % maxcol, j // maxcol of a grid of block_size cells,
    centred by calculate_start_point) one <path> per point -- "M prev L cur" while the pen is down,
    a bare "M cur" right after a pen lift -- stroked and filled in a colour that is re-drawn from
    random.randint(0, 225)^3 after every lift (utils.py:45,69), or "#000" when color_mode is False."""
    num_char = len(data)
    if num_char < 1:
        return None
    max_color_intensity = 225

    def new_color():
        return "rgb(%d,%d,%d)" % (random.randint(0, max_color_intensity), random.randint(0, max_color_intensity),
                                  random.randint(0, max_color_intensity))

    numrow = np.ceil(float(num_char) / float(maxcol))
    width, height = block_size * (min(num_char, maxcol)), block_size * numrow
    out = ['<?xml version="1.0" encoding="utf-8" ?>',
           '<svg baseProfile="full" height="%s" version="1.1" width="%s" xmlns="http://www.w3.org/2000/svg" '
           'xmlns:ev="http://www.w3.org/2001/xml-events" xmlns:xlink="http://www.w3.org/1999/xlink"><defs />'
           % (height, width),
           '<rect fill="white" height="%s" width="%s" x="0" y="0" />' % (height, width)]
    the_color = new_color()
    for j in range(num_char):
        lift_pen = 0
        cdata = data[j]
        abs_x, abs_y, _, _ = calculate_start_point(cdata, factor, block_size)
        abs_x += (j % maxcol) * block_size
        abs_y += (j // maxcol) * block_size
        for i in range(len(cdata)):
            x = round(float(cdata[i, 0]) * factor, 3)
            y = round(float(cdata[i, 1]) * factor, 3)
            prev_x, prev_y = round(abs_x, 3), round(abs_y, 3)
            abs_x += x
            abs_y += y
            if lift_pen == 1:
                p = "M %s,%s " % (abs_x, abs_y)
                the_color = new_color()
            else:
                p = "M %s,%s L %s,%s " % (prev_x, prev_y, abs_x, abs_y)
            lift_pen = max(cdata[i, 2], cdata[i, 3])  # lift pen on eos or eoc
            if not color_mode:
                the_color = "#000"
            out.append('<path d="%s" fill="%s" stroke="%s" stroke-width="%s" />' % (p, the_color, the_color, stroke_width))
    out.append("</svg>")
    text = "".join(out)
    if svg_filename:
        with open(svg_filename, "w") as f:
            f.write(text)
    return text


def draw_stroke_color(data, factor=1, svg_filename="sample.svg", stroke_width=1, block_size=200, maxcol=5,
                      svg_only=False, color_mode=True):
    """Reference utils.py:85-103: split one long stroke sequence at the eoc points and draw the pieces."""
    def split_sketch(d):  # one piece per eoc; the trailing stub without an eoc is dropped
        counter, result = 0, []
        for i in range(len(d)):
            if d[i, 3] > 0:
                result.append(d[counter:i + 1])
                counter = i + 1
        return result
    data = np.array(data, dtype=np.float32)
    return draw_stroke_color_array(split_sketch(data), factor, svg_filename, stroke_width, block_size, maxcol,
                                   svg_only, color_mode)


class SketchLoader(object):
    """The reference's data path (utils.py:105-283) with the per-point Python loop replaced by a
    per-sketch copy: batches are identical, element for element, for the same `np.random` state.

    What the reference's `next_seq` does, restated: a row of T+1 points walks the corpus from the
    current pointer; sketch i contributes its first max(1, len_i - 1) raw points (the last raw point is
    never emitted), the last contributed point is forced to (eos, eoc, cont) = (0, 1, 0), every other
    point gets cont = 1 unless its eos/eoc is set; a fully consumed sketch advances the pointer, and the
    pointer advances once more at the end of the row (utils.py:245-253).  Per row two `np.random.rand()`
    draws scale x and y by U[0.7, 1.3) (utils.py:235-236) and they are drawn before the walk.
    `data/<name>.cpkl` is a pickled list of float32 [N_i,4] arrays (dx, dy, eos, eoc); a Python 2 pickle
    loads with encoding='latin1'."""

    def __init__(self, batch_size=50, seq_length=300, scale_factor=1.0, data_filename="kanji", data_dir="./data"):
        self.data_dir = data_dir
        self.batch_size = batch_size
        self.seq_length = seq_length
        self.scale_factor = scale_factor  # raw coordinates are divided by this
        data_file = os.path.join(self.data_dir, data_filename + ".cpkl")
        if not os.path.exists(data_file):
            raw_data_dir = os.path.join(self.data_dir, data_filename)
            if not os.path.isdir(raw_data_dir):
                raise IOError("%s not found, and no raw SVG directory %s to build it from" % (data_file, raw_data_dir))
            print("creating training data cpkl file from raw source")
            self.length_data = self.preprocess(raw_data_dir, data_file)
        self.load_preprocessed(data_file)
        self.num_samples = len(self.raw_data)
        self.index = list(range(self.num_samples))
        self.reset_index_pointer()

    @staticmethod
    def svg_to_strokes(svgfile, len_data=None, line_length_threshold=10.0, min_points_per_path=1, max_points_per_path=3):
        """One SVG file -> float32 [N,4] = (dx, dy, eos, eoc), the reference's build_lines (utils.py:148-192):
        every `d` attribute in document order; a line contributes its end point, a cubic Bezier
        n = clamp(int(length / 10) + 1, 1, 3) points at t = 1/n .. 1 (the chord length decides n, utils.py:165-180);
        a path's first point is a pen-up move; the last point of a path ends the stroke (eos); coordinates become
        moves relative to the previous point, the last row ends the content (eoc), the first row is dropped.
        Segments that are neither lines nor cubics are reported and skipped, as the reference does."""
        import xml.etree.ElementTree as ET
        from . import svgpath
        rows = []
        for elem in ET.parse(svgfile).iter():
            if "d" not in elem.attrib:
                continue
            segs = svgpath.segments(elem.attrib["d"])
            for i, seg in enumerate(segs):
                if seg[0] not in ("L", "C"):
                    print("encountered an element that is not just a line or bezier ")
                    print("type: ", seg[0])
                    print(seg)
                    continue
                p0, p3 = seg[1], seg[-1]
                length = float(np.sqrt((p3.real - p0.real) ** 2 + (p3.imag - p0.imag) ** 2))
                if len_data is not None:
                    len_data.append(length)
                if seg[0] == "C":
                    n = min(max(int(length / line_length_threshold) + 1, min_points_per_path), max_points_per_path)
                    p1, p2 = seg[2], seg[3]
                    pts = []
                    for k in range(n + 1):
                        t = float(k) / float(n)
                        z = (1. - t) ** 3 * p0 + 3. * t * (1. - t) ** 2 * p1 + 3.0 * t ** 2 * (1.0 - t) * p2 + t ** 3 * p3
                        pts.append((z.real, z.imag))
                else:
                    pts = [(p0.real, p0.imag), (p3.real, p3.imag)]
                if i == 0:
                    rows.append([pts[0][0], pts[0][1], 0, 0])
                for k in range(1, len(pts)):
                    rows.append([pts[k][0], pts[k][1], 1 if (k == len(pts) - 1 and i == len(segs) - 1) else 0, 0])
        lines = np.array(rows, dtype=np.float32)
        lines[1:, 0:2] -= lines[0:-1, 0:2].copy()
        lines[-1, 3] = 1
        return lines[1:]

    def preprocess(self, data_dir, data_file):
        """data/<name>/**/*.svg -> data/<name>.cpkl (reference utils.py:124-217): a protocol-2 pickle of the list of
        per-file stroke arrays.  Files are visited in sorted order (the reference takes os.walk's order, which the
        file system decides; the order is reshuffled every epoch anyway).  Returns the segment lengths seen."""
        len_data = []
        files = []
        for dirpath, dirnames, filenames in os.walk(data_dir):
            dirnames.sort()
            files += [os.path.join(dirpath, f) for f in sorted(filenames) if f[-3:] == "svg"]
        sketch = []
        for f in files:
            print("processing " + f)
            sketch.append(self.svg_to_strokes(f, len_data))
        with open(data_file, "wb") as fh:
            pickle.dump(sketch, fh, protocol=2)
        return len_data

    def load_preprocessed(self, data_file):
        with open(data_file, "rb") as f:
            raw = pickle.load(f, encoding="latin1")
        self.raw_data = []
        self._emit = []  # per sketch: the [m,5] block a row copies while walking over it
        for d in raw:
            d = np.array(d, dtype=np.float32)
            d[:, 0:2] /= self.scale_factor
            self.raw_data.append(d)
            m = max(1, len(d) - 1)
            blk = np.zeros((m, 5), dtype=np.float32)
            blk[:, 0:4] = d[:m]
            blk[:, 4] = ((blk[:, 2] <= 0) & (blk[:, 3] <= 0)).astype(np.float32)
            blk[m - 1, 2:5] = (0.0, 1.0, 0.0)  # end of content overrides end of stroke
            self._emit.append(blk)

    def _next_row(self, n):
        sx = np.random.rand() * 0.6 + 0.7
        sy = np.random.rand() * 0.6 + 0.7
        row = np.empty((n, 5), dtype=np.float32)
        filled = 0
        while filled < n:
            blk = self._emit[self.index[self.pointer]]
            take = min(len(blk), n - filled)
            row[filled:filled + take] = blk[:take]
            filled += take
            if take == len(blk):
                self.tick_index_pointer()
        self.tick_index_pointer()
        row[:, 0] *= sx
        row[:, 1] *= sy
        return row

    def next_batch(self):
        """(inputs, targets) = (batch[:, :-1], batch[:, 1:]) of a [B, T+1, 5] batch (utils.py:258-268)."""
        batch = np.stack([self._next_row(self.seq_length + 1) for _ in range(self.batch_size)]).astype(np.float32)
        return batch[:, 0:-1], batch[:, 1:]

    def current_data(self):
        return self.raw_data[self.index[self.pointer]]

    def tick_index_pointer(self):
        self.pointer += 1
        if self.pointer >= len(self.raw_data):
            self.pointer = 0
            self.epoch_finished = True

    def reset_index_pointer(self):
        """New epoch: pointer to 0 and a fresh np.random.permutation of the sketch order (utils.py:279-283)."""
        self.pointer = 0
        self.epoch_finished = False
        self.index = np.random.permutation(self.index)


def make_synthetic_dataset(path, num_sketches=200, seed=0):
    """Writes a stand-in corpus in the exact kanji.cpkl format (protocol-2 pickle of a list of float32
    [N_i,4] arrays, utils.py:190-195,214-215): the real KanjiVG blob is not distributed with the reference."""
    rng = np.random.RandomState(seed)
    sketches = []
    for _ in range(num_sketches):
        n = int(rng.randint(20, 121))
        d = np.zeros((n, 4), dtype=np.float32)
        d[:, 0:2] = rng.normal(0.0, 8.0, size=(n, 2))
        d[:, 2] = rng.rand(n) < 0.12
        d[-1, 2] = 1
        d[-1, 3] = 1
        sketches.append(d)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        pickle.dump(sketches, f, protocol=2)
    return path


def default_init(handle, seed=0):
    """Flat fp32 parameter vector from TF 0.x's default initialiser (uniform_unit_scaling, +-sqrt(3/fan_in);
    _linear biases zero, GRU gate bias 1.0) for the tensors of `handle` -- what tf.initialize_all_variables()
    gives the reference graph (model.py:46-47 and the cells of model.py:14-27)."""
    import math
    rng = np.random.RandomState(seed)
    flat = np.zeros(handle.P, dtype=np.float32)
    for name, off, shape in handle.tensors():
        n = int(np.prod(shape))
        if name.endswith("/Bias"):
            v = np.full(n, 1.0 if "Gates" in name else 0.0)
        else:
            fan_in = float(np.prod(shape[:-1])) if len(shape) > 1 else 1.0
            lim = math.sqrt(3.0 / fan_in)
            v = rng.uniform(-lim, lim, size=n)
        flat[off:off + n] = v
    return flat


def synthetic_batch(B, T, seed=0, p_eos=0.10, p_eoc=0.015):
    """Seeded stand-in for SketchLoader.next_batch output: offsets ~ N(0, 0.5^2), pen state one-hot
    with P(eos, eoc, cont) = (0.10, 0.015, 0.885), targets = inputs shifted by one point.  Returns x, y [B,T,5]."""
    rng = np.random.RandomState(seed)
    seq = np.zeros((B, T + 1, 5), dtype=np.float64)
    seq[:, :, 0:2] = rng.normal(0.0, 0.5, size=(B, T + 1, 2))
    u = rng.rand(B, T + 1)
    pen = np.where(u < p_eoc, 1, np.where(u < p_eoc + p_eos, 0, 2))
    for k in range(3):
        seq[:, :, 2 + k] = pen == k
    seq = seq.astype(np.float32)
    return np.ascontiguousarray(seq[:, :-1]), np.ascontiguousarray(seq[:, 1:])



def synthetic_dropout_mask(B, T, H, keep_prob, seed=0):
    """Seeded 0/1 output-dropout mask [B,T,H] (what TF's floor(keep + U[0,1)) would be for a fixed draw,
    model.py:29-30) for runs that must be reproducible outside the library (bench.py's loss check)."""
    rng = np.random.RandomState(seed)
    return (rng.rand(B, T, H) < keep_prob).astype(np.float32)


class PrefetchingLoader(object):
    """SketchLoader.next_batch() (reference utils.py:258-268) produced one batch ahead by a background thread into
    page-locked buffers, so that train.py's timed region (train.py:83-87: next_batch + the step) is the step alone
    and skrnn_train_step_host copies straight from the batch's own memory.

    Same batches, same order, same `np.random` stream as the wrapped loader used directly: the worker is the only
    consumer of np.random while an epoch runs, it stops at the batch that finishes the epoch, and
    reset_index_pointer() (which draws the new permutation, utils.py:279-283) runs on the caller's thread while
    the worker is parked.  `pointer` / `epoch_finished` are those of the last batch handed out, which is what
    train.py prints and tests."""

    def __init__(self, loader, slots=3):
        import queue
        import threading
        self.loader = loader
        self.num_samples = loader.num_samples
        self.batch_size, self.seq_length = loader.batch_size, loader.seq_length
        self.pointer, self.epoch_finished = loader.pointer, loader.epoch_finished
        shape = (loader.batch_size, loader.seq_length, 5)
        self._bufs = [(self._alloc(shape), self._alloc(shape)) for _ in range(slots)]
        self._free = queue.Queue()
        self._ready = queue.Queue()
        for i in range(slots):
            self._free.put(i)
        self._held = None
        self._go = threading.Event()
        self._stop = False
        self._thread = threading.Thread(target=self._work, daemon=True)
        self._thread.start()
        self._started = False  # the worker only starts with the first next_batch() / reset_index_pointer() of the caller

    @staticmethod
    def _alloc(shape):
        try:
            import torch
            if torch.cuda.is_available():
                return torch.empty(shape, dtype=torch.float32).pin_memory().numpy()
        except Exception:
            pass
        return np.empty(shape, dtype=np.float32)

    def _work(self):
        while True:
            self._go.wait()
            if self._stop:
                return
            slot = self._free.get()
            if slot is None:
                return
            x, y = self.loader.next_batch()
            bx, by = self._bufs[slot]
            bx[...] = x
            by[...] = y
            done = self.loader.epoch_finished
            if done:
                self._go.clear()  # park until the caller starts the next epoch
            self._ready.put((slot, self.loader.pointer, done))

    def next_batch(self):
        if not self._started:
            self._started = True
            self._go.set()
        if self._held is not None:
            self._free.put(self._held)  # the previous batch's buffers may be overwritten now
        slot, pointer, done = self._ready.get()
        self._held = slot
        self.pointer, self.epoch_finished = pointer, done
        return self._bufs[slot]

    def reset_index_pointer(self):
        """Start of an epoch (train.py:79).  Only valid when the previous epoch was consumed to its end (or nothing
        was consumed yet), like the reference's loop does."""
        if self._started:
            # the worker may be mid-epoch: consume up to the batch that ends the epoch, where it parks
            while not self.epoch_finished:
                self.next_batch()
        if self._held is not None:
            self._free.put(self._held)
            self._held = None
        self.loader.reset_index_pointer()
        self.pointer, self.epoch_finished = self.loader.pointer, self.loader.epoch_finished
        self._started = True
        self._go.set()

    def close(self):
        self._stop = True
        self._go.set()
        self._free.put(None)
