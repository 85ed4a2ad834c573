#!/usr/bin/env python
"""train.py -- the reference's training driver (reference train.py) on the B200 library: same flags,
same log line, same save/<dataset>/ layout; the sess.run step (train.py:86) is one call into libskrnn."""
import argparse
import os
import time

from sketch_rnn_b200.model import Model, Saver, Session, save_config
from sketch_rnn_b200.utils import PrefetchingLoader, SketchLoader


def main():
    parser = argparse.ArgumentParser()
    parser.add_argument('--rnn_size', type=int, default=256, help='size of RNN hidden state')
    parser.add_argument('--num_layers', type=int, default=2, help='number of layers in the RNN')
    parser.add_argument('--model', type=str, default='lstm', help='rnn, gru, or lstm')
    parser.add_argument('--batch_size', type=int, default=100, help='minibatch size')
    parser.add_argument('--seq_length', type=int, default=300, help='RNN sequence length')
    parser.add_argument('--num_epochs', type=int, default=500, help='number of epochs')
    parser.add_argument('--save_every', type=int, default=250, help='save frequency')
    parser.add_argument('--grad_clip', type=float, default=5.0, help='clip gradients at this value')
    parser.add_argument('--learning_rate', type=float, default=0.005, help='learning rate')
    parser.add_argument('--decay_rate', type=float, default=0.99, help='decay rate for rmsprop')
    parser.add_argument('--num_mixture', type=int, default=24, help='number of gaussian mixtures')
    parser.add_argument('--data_scale', type=float, default=15.0, help='factor to scale raw data down by')
    parser.add_argument('--keep_prob', type=float, default=0.8, help='dropout keep probability')
    parser.add_argument('--stroke_importance_factor', type=float, default=200.0,
                        help='relative importance of pen status over mdn coordinate accuracy')
    parser.add_argument('--dataset_name', type=str, default="kanji", help='name of directory containing training data')
    # additions (not in the reference): arithmetic path and an early stop for smoke runs
    parser.add_argument('--precision', type=str, default="bf16", choices=["bf16", "f32"],
                        help='bf16 tensor-core path or fp32 reference-precision path')
    parser.add_argument('--max_batches', type=int, default=0, help='stop after this many batches (0 = run all epochs)')
    parser.add_argument('--no_prefetch', action='store_true',
                        help='build every batch inside the timed region, like the reference (default: one batch ahead '
                             'on a background thread, into page-locked buffers; same batches either way)')
    args = parser.parse_args()
    train(args)


def train(args):
    data_loader = SketchLoader(args.batch_size, args.seq_length, args.data_scale, args.dataset_name)
    if not args.no_prefetch:
        data_loader = PrefetchingLoader(data_loader)

    dirname = os.path.join('save', args.dataset_name)
    if not os.path.exists(dirname):
        os.makedirs(dirname)
    save_config(os.path.join(dirname, 'config.pkl'), args)

    model = Model(args)
    b_processed = 0
    batch_times = []  # addition: the log line keeps the reference's two decimals, the run's mean is printed at the end

    with Session(model) as sess:
        saver = Saver(model)

        def save_model():
            checkpoint_path = os.path.join('save', args.dataset_name, 'model.ckpt')
            saver.save(sess, checkpoint_path, global_step=b_processed)
            print("model saved to {}".format(checkpoint_path))

        for e in range(args.num_epochs):
            sess.run(model.assign_lr(args.learning_rate * (args.decay_rate ** e)))
            data_loader.reset_index_pointer()
            state = sess.run(model.initial_state)
            while not data_loader.epoch_finished:
                start = time.time()
                input_data, target_data = data_loader.next_batch()
                feed = {model.input_data: input_data, model.target_data: target_data, model.initial_state: state}
                train_loss, shape_loss, pen_loss, state, _ = sess.run(
                    [model.cost, model.cost_shape, model.cost_pen, model.final_state, model.train_op], feed)
                end = time.time()
                batch_times.append(end - start)
                b_processed += 1
                print("{}/{} (epoch {} batch {}), cost = {:.2f} ({:.2f}+{:.4f}), time/batch = {:.2f}"
                      .format(data_loader.pointer + e * data_loader.num_samples,
                              args.num_epochs * data_loader.num_samples,
                              e, b_processed, train_loss, shape_loss, pen_loss, end - start))
                assert train_loss < 30000  # if dodgy loss (incl. NaN), exit w/ error.
                if b_processed % args.save_every == 0 and b_processed > 0:
                    save_model()
                if args.max_batches and b_processed >= args.max_batches:
                    save_model()
                    report_time(args, batch_times)
                    return
        save_model()
        report_time(args, batch_times)


def report_time(args, batch_times):
    """Mean of the reference's timed region (next_batch + the step, train.py:83-87) after three warm-up batches."""
    steady = batch_times[3:] if len(batch_times) > 6 else batch_times
    if steady:
        ms = 1e3 * sum(steady) / len(steady)
        print("mean time/batch = {:.3f} ms over {} batches ({:.0f} stroke-steps/s)".format(
            ms, len(steady), args.batch_size * args.seq_length / (ms * 1e-3)))


if __name__ == '__main__':
    main()
