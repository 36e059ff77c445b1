"""Kernel-time breakdown of one real training step with torch.profiler (analysis only; numbers
reported by bench.py never come from a profiled run)."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import harness  # noqa: E402
from can_b200.model import CANNet  # noqa: E402
from can_b200.utils.train_eval_utils import FusedTrainer  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict("he"))
    m = m.cuda()
    x = harness.make_input(B, 768, 1024).cuda()
    gt = harness.make_gt(B, 768, 1024).cuda()
    tr = FusedTrainer(m, lr=1e-10)
    for _ in range(3):
        tr.step(x, gt)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            tr.step(x, gt)
        torch.cuda.synchronize()
    ev = [e for e in prof.key_averages() if e.device_time_total > 0]
    ev.sort(key=lambda e: -e.device_time_total)
    tot = sum(e.device_time_total for e in ev)
    print(f"device time per step: {tot / 3 / 1e3:.3f} ms over {sum(e.count for e in ev) // 3} kernels")
    for e in ev[:32]:
        print(f"{e.device_time_total / 3:9.1f} us {e.count // 3:4d}  {e.key[:90]}")
    cpu = [e for e in prof.key_averages() if e.self_cpu_time_total > 0]
    print(f"cpu self time per step: {sum(e.self_cpu_time_total for e in cpu) / 3 / 1e3:.2f} ms")


if __name__ == "__main__":
    main()
