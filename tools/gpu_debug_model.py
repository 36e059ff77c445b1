"""GPU bring-up probe: whole CANNet forward against the committed golden vectors."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import harness  # noqa: E402
from can_b200 import engine, ops  # noqa: E402
from can_b200.model import CANNet  # noqa: E402

CASES = {"he_1x64x96": ("he", 1, 64, 96), "he_2x48x80": ("he", 2, 48, 80), "he_1x104x136": ("he", 1, 104, 136),
         "default_1x64x64": ("default", 1, 64, 64), "he_1x128x192": ("he", 1, 128, 192)}


def main():
    golden = np.load(os.path.join(ROOT, "tests", "golden", "cannet_golden.npz"))
    precision = sys.argv[1] if len(sys.argv) > 1 else "bf16"
    for name, (regime, n, h, w) in CASES.items():
        m = CANNet(load_weights=True)
        m.load_state_dict(harness.make_state_dict(regime))
        m = m.cuda()
        m.precision = precision
        x = harness.make_input(n, h, w).cuda()
        with torch.no_grad():
            out, saved = engine.forward(m, x, m.nsplit, save=True)
        torch.cuda.synchronize()
        ref = torch.from_numpy(golden[f"{name}/f64/out"]).cuda()
        res = {"case": name, "precision": precision, "rel_l2": harness.rel_l2(out, ref),
               "rel_count": harness.rel_count(out, ref)}
        if f"{name}/fv" in golden.files:
            fv = saved["cat"].view(512, 0).to_nchw()
            fi = saved["cat"].view(512, 512).to_nchw()
            res["fv_rel"] = harness.rel_l2(fv, torch.from_numpy(golden[f"{name}/fv"]).cuda())
            res["fi_rel"] = harness.rel_l2(fi, torch.from_numpy(golden[f"{name}/fi"]).cuda())
        print("MODEL " + json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
