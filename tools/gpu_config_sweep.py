"""Throughput of the BASELINE.json configurations that are not the bench headline (CUDA events,
3 warm-up + 10 timed iterations each, inputs resident).  Output -> profiles/."""
import json
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

FWD_GFLOP = {(768, 1024): 733.36, (512, 512): 244.47, (2048, 2048): 3911.14, (384, 512): 183.36}

CONFIGS = [
    # name, train, batch, H, W, precision
    ("B infer 768x1024 b1 bf16", False, 1, 768, 1024, "bf16"),
    ("B infer 768x1024 b16 bf16", False, 16, 768, 1024, "bf16"),
    ("B infer 768x1024 b1 fp32", False, 1, 768, 1024, "fp32"),
    ("B infer 768x1024 b16 fp32", False, 16, 768, 1024, "fp32"),
    ("C train 512x512 b8 bf16", True, 8, 512, 512, "bf16"),
    ("D train 768x1024 b4 bf16", True, 4, 768, 1024, "bf16"),
    ("D train 768x1024 b4 fp32", True, 4, 768, 1024, "fp32"),
    ("E infer 2048x2048 b1 bf16", False, 1, 2048, 2048, "bf16"),
    ("E infer 2048x2048 b4 bf16", False, 4, 2048, 2048, "bf16"),
    ("E infer 2048x2048 b8 bf16", False, 8, 2048, 2048, "bf16"),
    ("A infer 384x512 b1 fp32", False, 1, 384, 512, "fp32"),
]


def main():
    sd = harness.make_state_dict("he")
    only = sys.argv[1:]
    for name, train, B, H, W, prec in CONFIGS:
        if only and not any(o in name for o in only):
            continue
        m = CANNet(load_weights=True)
        m.load_state_dict(sd)
        m = m.cuda()
        m.precision = prec
        x = harness.make_input(B, H, W).cuda()
        gt = harness.make_gt(B, H, W).cuda()
        tr = FusedTrainer(m, lr=1e-10) if train else None

        def step():
            if train:
                return tr.step(x, gt)
            with torch.no_grad():
                return ops.mse_count(m(x), gt, want_grad=False)[1]

        for _ in range(3):
            r = step()
        torch.cuda.synchronize()
        assert torch.isfinite(r).all(), name
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        ips = B / (ms * 1e-3)
        tf = ips * FWD_GFLOP[(H, W)] * (3 if train else 1) / 1e3
        print("CONFIG " + json.dumps({"config": name, "ms_per_step": round(ms, 3), "img_per_s": round(ips, 1),
                                      "model_tflops": round(tf, 1), "mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}),
              flush=True)
        del m, tr, x, gt
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()


if __name__ == "__main__":
    main()
