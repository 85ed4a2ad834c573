"""Data-parallel training and sharded sampling over the GPUs of one box (new relative to the reference,
which is single-process: SURVEY.md 8e).

One process per GPU.  Batch rows are independent within a step (own state, own eoc resets, own loss
terms; the loss is a mean over B*T rows), so rank r takes rows [row0, row0 + rows) of the global batch,
computes its share of the mean loss and of the gradient (`loss_rows` = B_global * T), the flat fp32
gradient is summed across ranks with one all-reduce, and every rank applies the identical clip + Adam.
Sampling shards sketches with no communication at all.
"""
import torch
import torch.distributed as dist


def shard_bounds(total, world, rank):
    """Contiguous, balanced split of `total` rows: returns (first_row, num_rows) of `rank`."""
    base, rem = divmod(total, world)
    rows = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, rows


def allreduce_sum_(tensors, group=None):
    """In-place sum over ranks of every tensor in `tensors` (no-op without an initialised group)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    for t in tensors:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)


def dp_train_step(handle, x_shard, y_shard, state, lr, seed, row0, global_rows, group=None):
    """One data-parallel step on this rank's shard.  Returns (losses[3] summed over ranks, new state)."""
    losses, state = handle.loss_grad(x_shard, y_shard, state, seed=seed, row0=row0, loss_rows=global_rows)
    allreduce_sum_([handle.grads_tensor(), losses], group)
    handle.apply_clip_adam(lr)
    return losses, state
