#!/usr/bin/env python
"""sample.py -- the reference's sampling driver (reference sample.py): same flags, loads
save/<dataset>/config.pkl and the latest checkpoint, draws sketches with the rejection filter of
sample_sketches (sample.py:69-103) and writes <filename>.svg."""
import argparse
import os
import sys

from sketch_rnn_b200.model import Model, Saver, Session, get_checkpoint_state, load_config
from sketch_rnn_b200.utils import calculate_start_point, draw_stroke_color_array


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('--filename', type=str, default='output', help='filename of .svg file to output, without .svg')
    parser.add_argument('--sample_length', type=int, default=600, help='number of strokes to sample')
    parser.add_argument('--picture_size', type=float, default=160, help='a centered svg will be generated of this size')
    parser.add_argument('--scale_factor', type=float, default=1, help='factor to scale down by for svg output.  smaller means bigger output')
    parser.add_argument('--num_picture', type=int, default=20, help='number of pictures to generate')
    parser.add_argument('--num_col', type=int, default=5, help='if num_picture > 1, how many pictures per row?')
    parser.add_argument('--dataset_name', type=str, default="kanji", help='name of directory containing training data')
    parser.add_argument('--color_mode', type=int, default=1, help='set to 0 if you are a black and white sort of person...')
    parser.add_argument('--stroke_width', type=float, default=2.0, help='thickness of pen lines')
    parser.add_argument('--temperature', type=float, default=0.1, help='sampling temperature')
    # addition (not in the reference): give up after this many rejected attempts
    parser.add_argument('--max_attempts', type=int, default=0, help='0 = keep trying like the reference')
    # addition: candidate sketches drawn per kernel launch (1 = the reference's one-by-one loop)
    parser.add_argument('--candidates_per_launch', type=int, default=64, help='sketches drawn per launch before filtering')
    return parser


def accept_sketch(strokes, min_size, max_size, min_num_stroke, max_num_stroke):
    """The filter of sample.py:85-95 on one sketch [n,5] (dx, dy already in data_scale units)."""
    _, _, num_stroke, num_char, _ = strokes.sum(0)
    if num_stroke < min_num_stroke or num_char == 0 or num_stroke > max_num_stroke:
        return False
    _, _, sizex, sizey = calculate_start_point(strokes)
    if sizex > max_size or sizey > max_size:
        return False
    if sizex < min_size or sizey < min_size:
        return False
    return True


def sample_sketches(model, sess, sample_args, min_size_ratio=0.0, max_size_ratio=0.8, min_num_stroke=4,
                    max_num_stroke=22, svg_only=True):
    """Rejection loop of sample.py:69-103: accept a sketch iff 4 <= #strokes <= 22, it contains an
    end-of-content point, and its bounding box fits in 0.8 * picture_size.

    Candidates are drawn `candidates_per_launch` at a time in one kernel launch (Model.sample_candidates) and
    filtered on the batch: the stroke / end-of-content counts for the whole batch in one NumPy pass (they reject
    most candidates), the rounded running-sum bounding box of utils.calculate_start_point only for the survivors.
    Candidates are examined in draw order and the first num_picture accepted ones are kept, so under the same
    np.random state the accepted sketches are the ones the reference's one-by-one loop accepts
    (candidates_per_launch = 1 is that loop)."""
    N = sample_args.num_picture
    frame_size = float(sample_args.picture_size)
    max_size = frame_size * max_size_ratio
    min_size = frame_size * min_size_ratio
    per_launch = max(1, int(getattr(sample_args, "candidates_per_launch", 1)))
    count, attempts = 0, 0
    sketch_list, param_list = [], []
    temp_mixture = temp_pen = sample_args.temperature
    while count < N:
        if sample_args.max_attempts and attempts >= sample_args.max_attempts:
            break
        n = per_launch
        if sample_args.max_attempts:
            n = min(n, sample_args.max_attempts - attempts)
        strokes, lengths, mixp = model.sample_candidates(n, sample_args.sample_length, temp_mixture, temp_pen,
                                                         stop_if_eoc=True)
        # rows past a sketch's length are zero, so the sums over the padded array are the sums over the sketch
        num_stroke, num_char = strokes[:, :, 2].sum(1), strokes[:, :, 3].sum(1)
        maybe = (num_stroke >= min_num_stroke) & (num_char != 0) & (num_stroke <= max_num_stroke)
        for k in range(n):
            attempts += 1
            sys.stdout.write('. ')
            if not maybe[k]:
                continue
            sk = strokes[k, :int(lengths[k])].copy()
            if not accept_sketch(sk, min_size, max_size, min_num_stroke, max_num_stroke):
                continue
            print(count + 1, "/", N)
            count += 1
            sketch_list.append(sk)
            param_list.append(model.mixture_params(mixp[k, :int(lengths[k])]))
            if count >= N:
                break
    draw_stroke_color_array(sketch_list, factor=sample_args.scale_factor, maxcol=sample_args.num_col,
                            svg_filename=sample_args.filename + '.svg', stroke_width=sample_args.stroke_width,
                            block_size=sample_args.picture_size, svg_only=svg_only,
                            color_mode=sample_args.color_mode != 0)
    return sketch_list, param_list


def main():
    sample_args = build_parser().parse_args()
    saved_args = load_config(os.path.join('save', sample_args.dataset_name, 'config.pkl'))
    model = Model(saved_args, True)
    sess = Session(model)
    saver = Saver(model)
    ckpt = get_checkpoint_state(os.path.join('save', sample_args.dataset_name))
    print("loading model: ", ckpt.model_checkpoint_path)
    saver.restore(sess, ckpt.model_checkpoint_path)
    print("Console mode")
    sample_sketches(model, sess, sample_args, svg_only=True)


if __name__ == '__main__':
    main()
