"""Data-parallel gradient exchange for the flat gradient arena (replaces the
DistributedDataParallel wrap at train.py:121-122 of the reference).

The arena is cut into buckets in the order backward produces them; each bucket's all-reduce is
issued asynchronously as soon as backward has enqueued the kernels that fill it, on whatever
backend the process group uses (NCCL over NVLink on the GPU box, gloo in the CPU tests).  The mean
over ranks (DDP semantics) is NOT applied here: it is folded into the fused SGD kernel as
``grad_scale = 1 / world_size``.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class GradBuckets:
    def __init__(self, arena: torch.Tensor, ranges: dict, group=None):
        """``ranges``: tag -> (begin, end) element offsets into ``arena`` (must tile it exactly)."""
        spans = sorted(ranges.values())
        assert spans[0][0] == 0 and spans[-1][1] == arena.numel(), "buckets must cover the arena"
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:])), "buckets must not overlap or leave gaps"
        self.arena, self.ranges, self.group = arena, dict(ranges), group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self._pending = []
        # diagnostic only (tools/gpu_scale8.sh): ranks run their steps without exchanging anything, which separates the
        # cost of the exchange from the cost of waiting for the slowest GPU
        self._nocomm = os.environ.get("CAN_B200_DP_NOCOMM", "0") == "1"

    def ready(self, tag: str):
        """Called when the gradients of bucket ``tag`` have been enqueued on the current stream."""
        if self.world > 1 and not self._nocomm:
            a, b = self.ranges[tag]
            self._pending.append(dist.all_reduce(self.arena[a:b], group=self.group, async_op=True))

    def reduce_scalar(self, t: torch.Tensor):
        if self.world > 1 and not self._nocomm:
            self._pending.append(dist.all_reduce(t, group=self.group, async_op=True))

    def wait(self):
        for w in self._pending:
            w.wait()
        self._pending = []

    @property
    def grad_scale(self) -> float:
        return 1.0 / self.world
