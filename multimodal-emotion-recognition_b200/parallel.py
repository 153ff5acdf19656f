"""Data parallelism by dialogue: one process per GPU, a full replica each, gradients all-reduced over
NCCL (NVLink 5 / NVSwitch) in buckets that are launched while the rest of the backward is still running.

The reference has no multi-GPU path at all (src/train.py:20 pins cuda:0); this is the one exchange step the hot path
needs.  The model's backward (one C call) reports finished gradient ranges of the flat arena through a callback
(`m2f_plan_backward_bucketed`); each range is all-reduced asynchronously on NCCL's own stream, which waits for exactly
the kernels enqueued so far, and the compute stream joins all of them at the end of the backward.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of `n_items` dialogues for `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class GradBucketReducer:
    """All-reduces ranges of a flat gradient tensor as they are reported.  Device-agnostic (gloo on CPU tensors in the
    tests, NCCL on the GPU): `on_bucket(lo, hi)` launches an async all-reduce of flat[lo:hi]; `finish()` makes the
    current stream (or the host, for gloo) wait for all of them and applies the 1/world averaging if requested."""

    def __init__(self, get_flat: Callable[[], torch.Tensor], group=None, average: bool = True):
        self.get_flat = get_flat
        self.group = group
        self.world = dist.get_world_size(group)
        self.average = average
        self.pending: List[Tuple[object, int, int]] = []
        self.log: List[Tuple[int, int]] = []  # ranges reduced in the last step (tests, bench reporting)
        self.bytes_last_step = 0

    def begin_step(self):
        self.log = []
        self.bytes_last_step = 0

    def on_bucket(self, lo: int, hi: int):
        if hi <= lo:
            return
        flat = self.get_flat()
        chunk = flat[lo:hi]
        if self.average and flat.is_cuda:
            work = dist.all_reduce(chunk, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        else:
            work = dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.pending.append((work, lo, hi))
        self.log.append((lo, hi))
        self.bytes_last_step += (hi - lo) * flat.element_size()

    def finish(self):
        flat = self.get_flat()
        for work, lo, hi in self.pending:
            work.wait()  # NCCL: the current stream waits (no host block); gloo: host wait
            if self.average and not flat.is_cuda:
                flat[lo:hi].div_(self.world)
        self.pending = []


class DataParallelM2FNet(torch.nn.Module):
    """Wraps a (CUDA) `M2FNet` for data-parallel training: broadcasts rank 0's parameters, then overlaps the bucketed
    gradient all-reduce with the backward.  `forward` is the wrapped model's; `train.train()` of the reference can be
    called with this object as `model` (it only uses __call__, train/eval, parameters, state_dict)."""

    def __init__(self, model, group=None, bucket_mb: float = 32.0, average: bool = True):
        super().__init__()
        self.module = model
        self.group = group
        flat = model.flat_parameters()
        dist.broadcast(flat, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        self.reducer = GradBucketReducer(model.flat_grads, group, average)
        model.set_grad_hooks(self._on_bucket, int(bucket_mb * (1 << 20) / 4), self.reducer.finish)
        self._first = True

    def _on_bucket(self, lo: int, hi: int):
        if self._first:
            self.reducer.begin_step()
            self._first = False
        self.reducer.on_bucket(lo, hi)

    def forward(self, text, audio, mask):
        self._first = True
        return self.module(text, audio, mask)

    def state_dict(self, *a, **k):
        return self.module.state_dict(*a, **k)

    def load_state_dict(self, *a, **k):
        return self.module.load_state_dict(*a, **k)
