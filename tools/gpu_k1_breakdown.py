"""Per-kernel times of the K1 pooling chain (torch.profiler, analysis only)."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
N, h, w = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (16, 96, 128)))
dense = len(sys.argv) > 4 and sys.argv[4] == "dense"
cat = ops.new_act(1, N, h, w, 512 if dense else 1024, dev, zero=True)
cat.buf.normal_()
fv = cat if dense else cat.view(512, 0)
w1 = torch.randn(4, 512, 512, device=dev) * 0.05
w2 = torch.randn(4, 512, 512, device=dev) * 0.05
for _ in range(3):
    ops.ctx_pool(fv, w1, w2)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10):
        ops.ctx_pool(fv, w1, w2)
    torch.cuda.synchronize()
for e in prof.key_averages():
    print(f"K1B {e.key[:70]:70s} n={e.count:3d} avg={e.device_time_total / max(e.count, 1):8.1f} us")
