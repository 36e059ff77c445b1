"""NVTX ranges around the sections of the step (the reference has no tracing; the new repo adds
ranges so an ncu / Nsight timeline of a training run names its parts).

Off unless ``CAN_B200_NVTX=1`` (read once at import): ``push`` / ``pop`` are then plain no-ops, so the step costs nothing
for them.  Ranges are host-side markers: inside a CUDA-graph capture they mark the capture, not the replay - the replay is
one range ("can.step.graph" / "can.dp.<bucket>") opened by the trainer around ``graph.replay()``.

Names: can.fwd.{frontend,context,backend}  can.bwd.{backend,context,frontend}  can.loss  can.sgd  can.dp.<bucket>
can.step  can.epoch.train  can.epoch.eval
"""
import os

import torch

ENABLED = os.environ.get("CAN_B200_NVTX", "0") == "1"

if ENABLED:
    def push(name: str) -> None:
        torch.cuda.nvtx.range_push(name)

    def pop() -> None:
        torch.cuda.nvtx.range_pop()
else:
    def push(name: str) -> None:
        return None

    def pop() -> None:
        return None


class span:
    """``with nvtx.span("can.step"):`` - closes the range on exceptions too."""
    __slots__ = ("name",)

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        push(self.name)
        return self

    def __exit__(self, *exc):
        pop()
        return False
