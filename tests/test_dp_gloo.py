"""world_size-2 gloo test (CPU) of the data-parallel contract: each rank computes its shard's share of
the mean loss / gradient (row0, loss_rows), one sum all-reduce, result == the single-process full batch.
The per-shard compute is the oracle here (no GPU in this test); the GPU kernels honour the same
(row0, loss_rows) contract, checked in tests/test_gpu_parity.py::test_sharded_loss_matches_full_batch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    from oracle import sketch_oracle as so
    from sketch_rnn_b200 import dp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    B, T = 7, 6  # 7 rows over 2 ranks: uneven shards 4 + 3
    cfg = so.Config(model="lstm", rnn_size=8, num_layers=2, num_mixture=3, batch_size=B, seq_length=T, keep_prob=0.8)
    params = so.init_params(cfg, seed=1, dtype=np.float64, trained_like=True)
    x, y = so.synthetic_batch(B, T, seed=2, dtype=np.float64, p_eoc=0.2)
    s0 = so.synthetic_state(cfg, B, dtype=np.float64)
    mask = so.synthetic_dropout_mask(cfg, B, T).astype(np.float64)
    first, rows = dp.shard_bounds(B, world, rank)
    sl = slice(first, first + rows)
    losses, grads, _, _ = so.loss_and_grads(cfg, params, x[sl], y[sl], [s[sl] for s in s0], mask[sl],
                                            loss_scale_rows=B * T)
    flat = torch.from_numpy(so.flatten(cfg, grads, np.float64))
    lt = torch.tensor(losses, dtype=torch.float64)
    # the segment-wise reduction the overlapped GPU step uses (output layer first, then the layers top down) must
    # give the same sums as one all-reduce of the whole vector
    flat2, lt2 = flat.clone(), lt.clone()
    spec = so.param_spec(cfg)
    sizes = [int(np.prod(s)) for _, s in spec]
    offs = np.concatenate([[0], np.cumsum(sizes)])
    head = int(offs[2])                                   # output_w, output_b; then (Matrix, Bias) per lstm layer
    segs = [(cfg.num_layers, 0, head)] + [(l, int(offs[2 + 2 * l]), int(offs[4 + 2 * l] - offs[2 + 2 * l]))
                                          for l in range(cfg.num_layers - 1, -1, -1)]
    assert sum(c for _, _, c in segs) == flat.numel()
    seen = []
    dp.reduce_segments_(flat2, segs, lt2, before_segment=seen.append)
    dp.allreduce_sum_([flat, lt])
    assert seen == [2, 1, 0] and torch.equal(flat, flat2) and torch.equal(lt, lt2)
    if rank == 0:
        full_l, full_g, _, _ = so.loss_and_grads(cfg, params, x, y, s0, mask)
        np.save(out, np.array([np.max(np.abs(flat.numpy() - so.flatten(cfg, full_g, np.float64))),
                               np.max(np.abs(lt.numpy() - np.array(full_l)))]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_sum_to_full_batch(tmp_path):
    out = str(tmp_path / "res.npy")
    mp.spawn(_worker, args=(2, 29517, out), nprocs=2, join=True)
    err = np.load(out)
    assert err[0] < 1e-12 and err[1] < 1e-12


def test_shard_bounds():
    from sketch_rnn_b200 import dp
    assert [dp.shard_bounds(800, 8, r) for r in (0, 7)] == [(0, 100), (700, 100)]
    parts = [dp.shard_bounds(7, 2, r) for r in range(2)]
    assert parts == [(0, 4), (4, 3)]
    assert sum(p[1] for p in [dp.shard_bounds(8192, 3, r) for r in range(3)]) == 8192
