"""Loss trajectory of FusedTrainer at the bench shape against the oracle's torch-CPU restatement."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import cannet_torch_cpu as T  # noqa: E402
from oracle import harness  # noqa: E402
from can_b200.model import CANNet  # noqa: E402
from can_b200.utils.train_eval_utils import FusedTrainer  # noqa: E402


def main():
    B, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    steps = int(sys.argv[4]) if len(sys.argv) > 4 else 4
    lr = float(sys.argv[5]) if len(sys.argv) > 5 else 1e-7
    sd = harness.make_state_dict("he")
    x, gt = harness.make_input(B, H, W), harness.make_gt(B, H, W)
    for prec in ("bf16", "fp32"):
        m = CANNet(load_weights=True)
        m.load_state_dict(sd)
        m = m.cuda()
        m.precision = prec
        tr = FusedTrainer(m, lr=lr)
        losses = [float(tr.step(x.cuda(), gt.cuda())) for _ in range(steps)]
        print(prec, ["%.6g" % v for v in losses], flush=True)
    torch.set_num_threads(os.cpu_count())
    ref_sd = {k: v.clone() for k, v in sd.items()}
    bufs = {}
    ref = [float(T.train_step(x, gt, ref_sd, bufs, lr, 0.95)) for _ in range(steps)]
    print("oracle fp32 cpu", ["%.6g" % v for v in ref], flush=True)


if __name__ == "__main__":
    main()
