"""GPU bring-up probe: parameter gradients of the whole step against the oracle (fp64 torch-CPU port)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import cannet_torch_cpu as T  # noqa: E402
from oracle import harness  # noqa: E402
from can_b200.model import CANNet  # noqa: E402


def main():
    precisions = sys.argv[1:] or ["bf16", "fp32"]
    for (regime, n, h, w) in [("he", 1, 64, 96), ("he", 2, 48, 80), ("default", 1, 104, 136)]:
        sd = harness.make_state_dict(regime)
        x = harness.make_input(n, h, w)
        gt = harness.make_gt(n, h, w)
        loss_ref, out_ref, gref = T.loss_and_grads(x, gt, sd, dtype=torch.float64)
        for prec in precisions:
            m = CANNet(load_weights=True)
            m.load_state_dict(sd)
            m = m.cuda()
            m.precision = prec
            out = m(x.cuda())
            loss = torch.nn.functional.mse_loss(out, gt.cuda(), reduction="sum")
            loss.backward()
            torch.cuda.synchronize()
            res = {"case": f"{regime}_{n}x{h}x{w}", "precision": prec, "out_rel": harness.rel_l2(out.detach().cpu(), out_ref),
                   "loss_rel": abs(float(loss) - float(loss_ref)) / abs(float(loss_ref))}
            worst = []
            for k, p in m.named_parameters():
                r = harness.rel_l2(p.grad.cpu(), gref[k])
                worst.append((r, k))
            worst.sort(reverse=True)
            res["worst"] = [(round(r, 6), k) for r, k in worst[:6]]
            res["median"] = round(sorted(r for r, _ in worst)[len(worst) // 2], 6)
            res["all"] = {k: round(r, 6) for r, k in worst}
            print("BWD " + json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
