"""Contextual module standalone (BASELINE config E / B shapes): K1 (multi-scale pooling + tables) and K2 (fused
contrast GEMM + sigmoid fusion, MODE_CTX2) timed with CUDA events against the HBM roofline.
Algorithmic bytes: K1 reads fv once (512*h*w*2 B per image, bf16); the whole module reads fv and writes fi
(2*512*h*w*2 B per image; the second read of fv by K2 is an L2 hit or not depending on size)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402


def timeit(fn, iters=20):
    """Device time of fn(): captured once into a CUDA graph and replayed (the real step runs inside a graph; eager
    timing of a 5-launch chain of ~10 us kernels would measure the Python / ctypes launch path instead)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    dev = torch.device("cuda:0")
    peak = 6540.8
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p)).get("hbm_gbs", peak)
    g = torch.Generator().manual_seed(0)
    w1 = torch.randn(4, 512, 512, generator=g).to(dev) * 0.05
    w2l = [(torch.randn(512, 512, 1, 1, generator=g) * 0.05).to(dev) for _ in range(4)]
    w2 = torch.stack([w.reshape(512, 512) for w in w2l])
    w2stack = ops.pack_ctx_w2(w2l, 1)
    for name, N, h, w in (("B 768x1024 b16", 16, 96, 128), ("E 2048x2048 b1", 1, 256, 256), ("E 2048x2048 b4", 4, 256, 256),
                          ("E 2048x2048 b8", 8, 256, 256)):
        cat = ops.new_act(1, N, h, w, 1024, dev, zero=True)
        cat.buf.normal_()
        fv, fi = cat.view(512, 0), cat.view(512, 512)
        wint = ops.ctx_wint(1, 1, h, w, dev)
        t1 = timeit(lambda: ops.ctx_pool(fv, w1, w2))
        pooled, tabA, tabB = ops.ctx_pool(fv, w1, w2)
        ttab = ops.ctx_ttab(tabA, tabB)
        t2 = timeit(lambda: ops.conv_tc(fv, w2stack, 2048, taps=1, mode=ops.MODE_CTX2, out=fi, ctx_wint=wint.buf, ctx_ttab=ttab))
        b1 = N * h * w * 512 * 2
        flop2 = 2.0 * 512 * 2048 * N * h * w
        print("CTX " + json.dumps({
            "config": name, "K1_us": round(t1 * 1e6, 1), "K1_GBps": round(b1 / t1 / 1e9, 1), "K1_frac_hbm": round(b1 / t1 / 1e9 / peak, 3),
            "K2_us": round(t2 * 1e6, 1), "K2_TFLOPs": round(flop2 / t2 / 1e12, 1),
            "module_us": round((t1 + t2) * 1e6, 1), "module_GBps_fv_in_fi_out": round(2 * b1 / (t1 + t2) / 1e9, 1),
            "module_frac_hbm": round(2 * b1 / (t1 + t2) / 1e9 / peak, 3)}), flush=True)
        del cat


if __name__ == "__main__":
    main()
