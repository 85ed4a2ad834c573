"""Data-parallel training and sharded sampling over the GPUs of one box (new relative to the reference,
which is single-process: SURVEY.md 8e).

One process per GPU.  Batch rows are independent within a step (own state, own eoc resets, own loss
terms; the loss is a mean over B*T rows), so rank r takes rows [row0, row0 + rows) of the global batch,
computes its share of the mean loss and of the gradient (`loss_rows` = B_global * T), the flat fp32
gradient is summed across ranks with one all-reduce, and every rank applies the identical clip + Adam.
Sampling shards sketches with no communication at all.
"""
import os

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


_comm_streams = {}


def _comm_stream(device):
    st = _comm_streams.get(device)
    if st is None:
        st = _comm_streams[device] = torch.cuda.Stream(device=device)
    return st


def allreduce_grads_overlapped_(handle, losses, group=None):
    """Sum of the flat gradient (and the three losses) over the ranks, started segment by segment while the backward
    pass is still running: skrnn_loss_grad finishes the output layer's gradient first and the cell layers from the top
    down, and records an event for each; a communication stream waits for a segment's event and reduces that slice
    (NCCL over NVLink), so only the last segment's reduction is exposed.  The caller's stream then waits for the
    communication stream.  Same sums as one all-reduce of the whole vector."""
    g = handle.grads_tensor()
    dev = g.device
    comm = _comm_stream(dev)

    def gate(idx):
        handle.stream_wait_grad(idx, comm)

    with torch.cuda.stream(comm):
        reduce_segments_(g, handle.grad_segments(), losses, group, before_segment=gate)
    torch.cuda.current_stream(dev).wait_stream(comm)


def reduce_segments_(g, segments, losses, group=None, before_segment=None):
    """The segment-wise sum behind allreduce_grads_overlapped_: `segments` = [(index, offset, count)] covering the flat
    gradient `g`; `before_segment(index)` is called before a segment's collective is enqueued (the GPU path makes the
    communication stream wait for that segment's event there).  The losses ride with the first segment."""
    first = True
    for idx, off, cnt in segments:
        if before_segment is not None:
            before_segment(idx)
        if first:
            dist.all_reduce(losses, op=dist.ReduceOp.SUM, group=group)
            first = False
        dist.all_reduce(g[off:off + cnt], op=dist.ReduceOp.SUM, group=group)


def dp_train_step(handle, x_shard, y_shard, state, lr, seed, row0, global_rows, group=None, timing=None, overlap=True):
    """One data-parallel step on this rank's shard.  Returns (losses[3] summed over ranks, new state).
    overlap: reduce the gradient per layer while the layers below are still in their backward pass
    (allreduce_grads_overlapped_); False: one all-reduce of the whole vector after the backward pass.
    timing: optional dict; when given, the all-reduce runs un-overlapped between CUDA events and its milliseconds are
    added to timing["allreduce_ms"] (this synchronises the stream -- measurement passes only)."""
    losses, state = handle.loss_grad(x_shard, y_shard, state, seed=seed, row0=row0, loss_rows=global_rows)
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    if timing is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if multi:
            # line the ranks up first, so that the interval below is the all-reduce and not the wait for the slowest rank
            dist.all_reduce(torch.zeros(1, device=losses.device), group=group)
        e0.record()
        allreduce_sum_([handle.grads_tensor(), losses], group)
        e1.record()
        e1.synchronize()
        timing["allreduce_ms"] = timing.get("allreduce_ms", 0.0) + e0.elapsed_time(e1)
    elif multi and overlap and losses.is_cuda and os.environ.get("SKRNN_DP_OVERLAP", "1") != "0":
        allreduce_grads_overlapped_(handle, losses, group)
    else:
        allreduce_sum_([handle.grads_tensor(), losses], group)
    handle.apply_clip_adam(lr)
    return losses, state


class HostStep(object):
    """The training step a data-parallel caller makes with HOST buffers, the N-GPU counterpart of
    skrnn_train_step_host (one sess.run([cost, cost_shape, cost_pen, final_state, train_op], feed), train.py:85-86):
    this rank's rows of the batch go from page-locked host memory to the GPU, loss_grad runs, the flat gradient is
    all-reduced (when there is more than one rank), clip + Adam, and the three losses and the final state come back
    to page-locked host memory.  Buffers are allocated once."""

    def __init__(self, handle, rows, T, global_rows, row0=0, group=None):
        dev = "cuda:%d" % handle.device
        self.h, self.rows, self.T, self.global_rows, self.row0, self.group = handle, rows, T, global_rows, row0, group
        self.x = torch.empty((rows, T, 5), dtype=torch.float32, device=dev)
        self.y = torch.empty_like(self.x)
        self.state = torch.zeros((handle.L, rows, handle.S), dtype=torch.float32, device=dev)
        self.losses_host = torch.empty(3, dtype=torch.float32).pin_memory()
        self.state_host = torch.zeros((handle.L, rows, handle.S), dtype=torch.float32).pin_memory()
        self.h2d_bytes = 2 * self.x.numel() * 4 + self.state.numel() * 4
        self.d2h_bytes = 12 + self.state.numel() * 4

    def __call__(self, x_pinned, y_pinned, lr, seed):
        """x_pinned, y_pinned: page-locked torch tensors [rows, T, 5].  Returns the losses (host, valid on return)."""
        self.x.copy_(x_pinned, non_blocking=True)
        self.y.copy_(y_pinned, non_blocking=True)
        self.state.copy_(self.state_host, non_blocking=True)
        losses, st = dp_train_step(self.h, self.x, self.y, self.state, lr, seed, self.row0, self.global_rows, self.group)
        self.losses_host.copy_(losses, non_blocking=True)
        self.state_host.copy_(st, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.losses_host
