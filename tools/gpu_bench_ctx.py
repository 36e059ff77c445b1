"""Micro-benchmark of the contextual-fusion kernel alone (CUDA events, batch 4, 96x128 feature map)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "ctx2"
    N, h, w = 4, 96, 128
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    cat = ops.new_act(1, N, h, w, 1024, dev, zero=True)
    cat.buf.copy_(torch.rand(1, N, h, w, 1024, generator=g).to(dev))
    fv, fi = cat.view(512, 0), cat.view(512, 512)
    w2 = [(torch.randn(512, 512, 1, 1, generator=g) * 0.05).to(dev) for _ in range(4)]
    w1 = [(torch.randn(512, 512, 1, 1, generator=g) * 0.05).to(dev) for _ in range(4)]
    w2stack = ops.pack_ctx_w2(w2, 1)
    pooled, tabA, tabB = ops.ctx_pool(fv, ops.stack_1x1(w1), ops.stack_1x1(w2))
    wint = ops.ctx_wint(1, 1, h, w, dev)
    ttab = ops.ctx_ttab(tabA, tabB)

    def run():
        if mode == "ctx2bwd":
            ops.conv_tc(fv, w2stack, 2048, taps=1, mode=ops.MODE_CTX2_BWD, out=run.D, mask=fi, ctx_wint=wint.buf, ctx_ttab=ttab)
        elif mode == "ctx2":
            ops.conv_tc(fv, w2stack, 2048, taps=1, mode=ops.MODE_CTX2, out=fi, ctx_wint=wint.buf, ctx_ttab=ttab)
        elif mode == "ctx":
            ops.conv_tc(fv, w2stack, 2048, taps=1, mode=ops.MODE_CTX, out=fi, tabA=tabA, tabB=tabB)
        else:  # plain GEMM of the same shape for reference
            out = run.out
            ops.conv_tc(fv, w2stack, 2048, taps=1, mode=ops.MODE_STORE, relu=False, out=out)
    run.out = ops.new_act(1, N, h, w, 2048, dev)
    run.D = ops.new_act(1, N, h, w, 4096, dev)
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
    print(f"{mode} dbg={os.environ.get('CAN_B200_DBG', '0')}: {ms * 1e3:.1f} us  ({2 * 512 * 2048 * N * h * w / ms / 1e9:.0f} TF)")


if __name__ == "__main__":
    main()
