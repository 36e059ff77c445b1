"""SM clock / power while one conv_tc configuration runs back to back (NVML samples from a side thread)."""
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402
import pynvml  # noqa: E402


def main():
    N, H, W, Cin, Cout = [int(v) for v in sys.argv[1:6]]
    secs = float(sys.argv[6]) if len(sys.argv) > 6 else 2.0
    dev = torch.device("cuda:0")
    x = ops.new_act(1, N, H, W, Cin, dev, zero=True)
    x.buf.normal_()
    w = torch.randn(Cout, Cin, 3, 3, device=dev) * 0.05
    wp = ops.pack_conv_weight(w, 1)
    b = torch.zeros(Cout, device=dev)
    out = ops.new_act(1, N, H, W, Cout, dev)
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    samples = []
    stop = False

    def sampler():
        while not stop:
            samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0))
            time.sleep(0.02)

    def run():
        ops.conv_tc(x, wp, Cout, mode=ops.MODE_STORE, relu=True, bias=b, out=out)

    for _ in range(5):
        run()
    torch.cuda.synchronize()
    th = threading.Thread(target=sampler)
    th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    n = 0
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(20):
            run()
        n += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    stop = True
    th.join()
    ms = e0.elapsed_time(e1) / n
    clk = sorted(s[0] for s in samples)
    pw = sorted(s[1] for s in samples)
    tag = " ".join(f"{k[9:]}={os.environ[k]}" for k in ("CAN_B200_DBG",) if k in os.environ)
    print(f"{Cin}->{Cout} {H}x{W} N={N} [{tag}]: {ms * 1e3:.1f} us  {2 * 9 * Cin * Cout * N * H * W / ms / 1e9:.0f} TF; "
          f"SM clock median {clk[len(clk) // 2]} MHz (min {clk[0]}), power median {pw[len(pw) // 2]:.0f} W (max {pw[-1]:.0f}), {len(samples)} samples")


if __name__ == "__main__":
    main()
