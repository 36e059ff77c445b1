import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops
dev = torch.device("cuda:0")
N, h, w = 16, 96, 128
cat = ops.new_act(1, N, h, w, 1024, dev, zero=True)
cat.buf.normal_()
w1 = torch.randn(4, 512, 512, device=dev) * 0.05
w2 = torch.randn(4, 512, 512, device=dev) * 0.05
for _ in range(3):
    ops.ctx_pool(cat.view(512, 0), w1, w2)
torch.cuda.synchronize()
print("ok")
