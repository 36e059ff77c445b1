"""Per-launch CUDA-event timing of the tensor-core kernels inside a real step (no profiler attached)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import harness  # noqa: E402
from can_b200 import ops  # noqa: E402
from can_b200.model import CANNet  # noqa: E402
from can_b200.utils.train_eval_utils import FusedTrainer  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    train = (sys.argv[2] if len(sys.argv) > 2 else "train") == "train"
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict("he"))
    m = m.cuda()
    x = harness.make_input(B, 768, 1024).cuda()
    gt = harness.make_gt(B, 768, 1024).cuda()
    tr = FusedTrainer(m, lr=1e-10) if train else None

    def step():
        if train:
            tr.step(x, gt)
        else:
            with torch.no_grad():
                m(x)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record()
    torch.cuda.synchronize()
    print(f"plain step: {e0.elapsed_time(e1) / 5:.3f} ms (batch {B})")
    import time
    if tr is not None:
        tr.use_graph = False
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    step()
    step()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"host enqueue time: {(t1 - t0) / 2 * 1e3:.3f} ms per step")
    ops.PROFILE = []
    os.environ["CAN_B200_GRAPH_INFER"] = "0"   # per-launch events need the eager wrappers (inference graph off too)
    if tr is not None:
        tr.use_graph = False
    step()
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    tot = 0.0
    for a, b, f, label in prof:
        ms = a.elapsed_time(b)
        tot += ms
        print(f"{ms * 1e3:9.1f} us {f / ms / 1e9:8.1f} TF  {label}")
    print(f"tensor kernels: {tot:.3f} ms, {sum(p[2] for p in prof) / tot / 1e9:.1f} TF avg")


if __name__ == "__main__":
    main()
