"""The GPU incumbent on the same box: the reference ships no
kernels, so what this repo has to beat on a B200 is the stock-op graph of model/CANNet.py:39-91 under torch eager
(cuDNN / cuBLAS / ATen) with the step body of utils/train_eval_utils.py:28-52 and torch.optim.SGD (train.py:125-126).

  python tools/gpu_incumbent.py                               # one GPU: fp32 (TF32 off / on), channels_last + bf16 autocast
  python -m torch.distributed.run --nproc-per-node 8 tools/gpu_incumbent.py --ddp   # stock DistributedDataParallel (train.py:121-122)

Workload = bench.py's train_768x1024_b4 (batch 4 per GPU of 3x768x1024, fwd + sum-MSE + bwd + SGD momentum 0.95), same
synthetic inputs and He-initialised weights, lr 1e-10 x world; 5 warm-up + 20 timed steps, CUDA events, max over ranks.
One JSON line per variant.  A tool, not part of the product path or of bench.py.
"""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import harness  # noqa: E402  (weight / input generators)


class StockCANNet(nn.Module):
    """model/CANNet.py's layer list and forward graph on stock torch.nn / torch.nn.functional ops (same 42 keys)."""

    def __init__(self):
        super().__init__()

        def seq(cfg, cin, dil):
            layers = []
            for v in cfg:
                if v == "M":
                    layers.append(nn.MaxPool2d(2, 2))
                else:
                    layers += [nn.Conv2d(cin, v, 3, padding=dil, dilation=dil), nn.ReLU(inplace=True)]
                    cin = v
            return nn.Sequential(*layers)

        self.frontend = seq([64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512], 3, 1)
        self.backend = seq([512, 512, 512, 256, 128, 64], 1024, 2)
        self.output_layer = nn.Conv2d(64, 1, 1)
        for s in (1, 2, 3, 6):
            setattr(self, f"conv{s}_1", nn.Conv2d(512, 512, 1, bias=False))
            setattr(self, f"conv{s}_2", nn.Conv2d(512, 512, 1, bias=False))

    def forward(self, x):
        fv = self.frontend(x)
        h, w = fv.shape[2], fv.shape[3]
        num, den = 0, 0
        for s in (1, 2, 3, 6):
            a = getattr(self, f"conv{s}_1")(F.adaptive_avg_pool2d(fv, (s, s)))
            sj = F.interpolate(a, size=(h, w), mode="bilinear", align_corners=True)
            wj = torch.sigmoid(getattr(self, f"conv{s}_2")(sj - fv))
            num = num + wj * sj
            den = den + wj
        fi = num / (den + 0.000000000001)
        return self.output_layer(self.backend(torch.cat((fv, fi), 1)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ddp", action="store_true")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--variants", default="fp32_tf32off,fp32_tf32on,bf16_autocast_channels_last")
    args = ap.parse_args()
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1")) if args.ddp else 1
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if args.ddp:
        dist.init_process_group("nccl", device_id=dev)
    B, H, W = args.batch, 768, 1024
    x = harness.make_input(B, H, W, seed=1234 + rank).to(dev)
    gt = harness.make_gt(B, H, W, seed=4321 + rank).to(dev)
    sd = harness.make_state_dict("he")
    crit = nn.MSELoss(reduction="sum")
    for variant in args.variants.split(","):
        tf32 = variant != "fp32_tf32off"
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True
        m = StockCANNet()
        m.load_state_dict(sd)
        m = m.to(dev)
        xin = x
        if variant.startswith("bf16"):
            m = m.to(memory_format=torch.channels_last)
            xin = x.contiguous(memory_format=torch.channels_last)
        net = nn.parallel.DistributedDataParallel(m, device_ids=[local]) if args.ddp else m
        opt = torch.optim.SGD(m.parameters(), lr=1e-10 * world, momentum=0.95, weight_decay=0)

        def step():
            opt.zero_grad()
            if variant.startswith("bf16"):
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    et = net(xin)
                loss = crit(et.float(), gt)
            else:
                loss = crit(net(xin), gt)
            loss.backward()
            opt.step()
            return loss

        for _ in range(args.warmup):
            step()
        torch.cuda.synchronize()
        if args.ddp:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            loss = step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if args.ddp:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        if rank == 0:
            print(json.dumps({"incumbent": "torch eager stock ops" + (" + DistributedDataParallel" if args.ddp else ""),
                              "variant": variant, "n_gpus": world, "per_gpu_batch": B, "H": H, "W": W,
                              "img_per_s": world * B * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
                              "steps": args.steps, "warmup": args.warmup, "final_loss": float(loss),
                              "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}), flush=True)
        del m, net, opt
        torch.cuda.empty_cache()
    if args.ddp:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
