"""128->128 data gradient (mask + bias gradient) at 4 x 384x512: staging-buffer knobs, one process (env read per launch)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402


def timeit(run):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 100


def main():
    dev = torch.device("cuda:0")
    for (N, H, W, Cin, Cout) in [(4, 384, 512, 128, 128), (4, 192, 256, 256, 256), (4, 96, 128, 512, 512)]:
        x = ops.new_act(1, N, H, W, Cin, dev, zero=True)
        x.buf.normal_()
        wp = ops.pack_conv_weight(torch.randn(Cout, Cin, 3, 3, device=dev) * 0.05, 1)
        out = ops.new_act(1, N, H, W, Cout, dev)
        mask = ops.new_act(1, N, H, W, Cout, dev, zero=True)
        mask.buf.normal_()
        db = torch.zeros(Cout, device=dev)
        b = torch.zeros(Cout, device=dev)
        cases = [("fwd store", dict(relu=True, bias=b)), ("dgrad mask", dict(relu=False, mask=mask)),
                 ("dgrad mask+db", dict(relu=False, mask=mask, db=db)), ("dgrad db only", dict(relu=False, db=db))]
        for env in [{}, {"CAN_B200_MASKBUFS": "0"}, {"CAN_B200_MASKBUFS": "2"}, {"CAN_B200_TMASTORE": "0"}]:
            for k in ("CAN_B200_MASKBUFS", "CAN_B200_TMASTORE"):
                os.environ.pop(k, None)
            os.environ.update(env)
            for name, kw in cases:
                us = timeit(lambda: ops.conv_tc(x, wp, Cout, mode=ops.MODE_STORE, out=out, **kw))
                print(f"{Cin}->{Cout} {H}x{W} N={N} {name:14s} {env}: {us:7.1f} us ({2 * 9 * Cin * Cout * N * H * W / us / 1e6:.0f} TF)",
                      flush=True)


if __name__ == "__main__":
    main()
