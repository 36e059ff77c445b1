"""Micro-benchmark of one conv_tc layer (CUDA events)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402


def main():
    N, H, W, Cin, Cout, dil = [int(v) for v in sys.argv[1:7]]
    mode = sys.argv[7] if len(sys.argv) > 7 else "store"
    dev = torch.device("cuda:0")
    x = ops.new_act(1, N, H, W, Cin, dev, zero=True)
    x.buf.normal_()
    w = torch.randn(Cout, Cin, 3, 3, device=dev) * 0.05
    wp = ops.pack_conv_weight(w, 1)
    b = torch.zeros(Cout, device=dev)
    if mode == "pool":
        out = ops.new_act(1, N, H // 2, W // 2, Cout, dev)
        arg = torch.empty(N, H // 2, W // 2, Cout, device=dev, dtype=torch.uint8)

        def run():
            ops.conv_tc(x, wp, Cout, dil=dil, mode=ops.MODE_POOL, bias=b, out=out, arg=arg)
    elif mode.startswith("dgrad"):
        out = ops.new_act(1, N, H, W, Cout, dev)
        mask = ops.new_act(1, N, H, W, Cout, dev, zero=True)
        mask.buf.normal_()
        db = torch.zeros(Cin, device=dev) if mode == "dgraddb" else None

        def run():
            ops.conv_tc(x, wp, Cout, dil=dil, mode=ops.MODE_STORE, relu=False, out=out, mask=mask, db=db)
    else:
        out = ops.new_act(1, N, H, W, Cout, dev)

        def run():
            ops.conv_tc(x, wp, Cout, dil=dil, mode=ops.MODE_STORE, relu=True, bias=b, out=out)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    tag = " ".join(f"{k[9:]}={os.environ[k]}" for k in ("CAN_B200_DBG", "CAN_B200_HALO", "CAN_B200_CTAS", "CAN_B200_STAGES",
                                                       "CAN_B200_BN", "CAN_B200_TMASTORE", "CAN_B200_PAIR") if k in os.environ)
    print(f"{Cin}->{Cout} {H}x{W} N={N} {mode} [{tag}]: {ms * 1e3:.1f} us  ({2 * 9 * Cin * Cout * N * H * W / ms / 1e9:.0f} TF)")


if __name__ == "__main__":
    main()
