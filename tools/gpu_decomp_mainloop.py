"""Where does the main loop of the 128- and 256-wide conv tiles spend its time?  One process, same binary; bring-up bits of
CAN_B200_DBG (read at every launch): 16 no epilogue | 1024 no A-tile loads | 2048 no B-tile loads | 4096 no MMAs (loads and
hand-shakes only); these three exist only in the CAN_DECOMP build (make decomp).  Results are garbage with any bit set; only the time is read.  CUDA events, 3 warm-up + 10 timed."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
# the bring-up build (`make -C can-distributed-pytorch_b200/csrc decomp`): the product library has no such switches in its loops
os.environ.setdefault("CAN_B200_LIB", os.path.join(ROOT, "can-distributed-pytorch_b200", "can_b200", "libcan_b200_decomp.so"))
if not os.path.exists(os.environ["CAN_B200_LIB"]):
    sys.exit(f"{os.environ['CAN_B200_LIB']} not found: build it with `make -C can-distributed-pytorch_b200/csrc decomp`")
from can_b200 import ops  # noqa: E402

SHAPES = [(4, 384, 512, 128, 128), (4, 96, 128, 512, 512), (4, 192, 256, 256, 256), (4, 192, 256, 128, 256)]
VARIANTS = [("1", "1", "pairs, halo A"), ("1", "0", "pairs, A per tap"), ("0", "0", "single CTA")]
BITS = [(0, "all"), (16, "no epilogue"), (16 + 1024, "no epi, no A loads"), (16 + 2048, "no epi, no B loads"),
        (16 + 1024 + 2048, "no epi, no loads"), (16 + 4096, "no epi, no MMAs"), (16 + 4096 + 1024, "no epi, no MMAs, no A"),
        (16 + 4096 + 2048, "no epi, no MMAs, no B")]


def main():
    dev = torch.device("cuda:0")
    for N, H, W, Cin, Cout in SHAPES:
        x = ops.new_act(1, N, H, W, Cin, dev, zero=True)
        x.buf.normal_()
        wp = ops.pack_conv_weight(torch.randn(Cout, Cin, 3, 3, device=dev) * 0.05, 1)
        b = torch.zeros(Cout, device=dev)
        out = ops.new_act(1, N, H, W, Cout, dev)
        for pair, haloa, vname in VARIANTS:
            os.environ["CAN_B200_PAIR"] = pair
            os.environ["CAN_B200_HALOA"] = haloa
            for stages in ("0",):
                os.environ["CAN_B200_STAGES"] = stages
                for dbg, what in BITS:
                    if stages == "4" and dbg not in (0, 16):
                        continue
                    os.environ["CAN_B200_DBG"] = str(dbg)

                    def run():
                        ops.conv_tc(x, wp, Cout, mode=ops.MODE_STORE, relu=True, bias=b, out=out)
                    for _ in range(3):
                        run()
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(10):
                        run()
                    e1.record()
                    torch.cuda.synchronize()
                    us = e0.elapsed_time(e1) * 100
                    print(f"{Cin}->{Cout} {H}x{W} N={N} [{vname:16s}] {what:28s}: {us:7.1f} us "
                          f"({2 * 9 * Cin * Cout * N * H * W / us / 1e6:.0f} TF)", flush=True)
    os.environ["CAN_B200_DBG"] = "0"


if __name__ == "__main__":
    main()
