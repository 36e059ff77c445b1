"""HBM-bound kernels against the measured copy bandwidth (CUDA events, inputs larger than the 126 MB L2)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from can_b200 import ops  # noqa: E402


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    dev = torch.device("cuda:0")
    peak = 6540.8
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p)).get("hbm_gbs", peak)
    out = []
    # K1: multi-scale pooling + tables, batch 16 of 96x128x512 bf16 inside the 1024-channel concat buffer
    N, h, w = 16, 96, 128
    cat = ops.new_act(1, N, h, w, 1024, dev, zero=True)
    cat.buf.normal_()
    fv = cat.view(512, 0)
    w1 = torch.randn(4, 512, 512, device=dev) * 0.05
    w2 = torch.randn(4, 512, 512, device=dev) * 0.05
    t = timeit(lambda: ops.ctx_pool(fv, w1, w2))
    b = N * h * w * 512 * 2
    out.append(dict(kernel="K1 ctx_pool_fwd (rows + finalize + 2 table GEMVs), batch 16", bytes=b, us=t * 1e6, GBps=b / t / 1e9))
    # loss + counts + grad, 64 images of 96x128 (two reads + one write)
    et = torch.randn(1024, 1, h, w, device=dev)
    gt = torch.rand(1024, 1, h, w, device=dev)
    t = timeit(lambda: ops.mse_count(et, gt, want_grad=True))
    b = 3 * et.numel() * 4
    out.append(dict(kernel="mse_count_fwd_bwd, 1024 x 96x128", bytes=b, us=t * 1e6, GBps=b / t / 1e9))
    # fused SGD over 4 parameter arenas' worth (> L2)
    n = 4 * 20_719_937
    pa, ga, ba = (torch.randn(n, device=dev) for _ in range(3))
    t = timeit(lambda: ops.sgd_momentum_step(pa, ga, ba, 1e-7, 0.95, 1.0, first=False))
    b = 5 * n * 4
    out.append(dict(kernel="sgd_momentum_step, 82.9 M floats", bytes=b, us=t * 1e6, GBps=b / t / 1e9))
    # data path: one 1536x2048 RGB sample (uint8 in, normalised fp32 NCHW + density map /8 out)
    img = torch.randint(0, 256, (1539, 2051, 3), dtype=torch.uint8, device=dev)
    dm = torch.rand(1539, 2051, device=dev)
    t = timeit(lambda: ops.crowd_sample(img, dm, flip=True))
    b = img.numel() + dm.numel() * 4 // 64 * 4 + 3 * 1536 * 2048 * 4 + 192 * 256 * 4   # dmap: 4 taps per output pixel
    out.append(dict(kernel="crowd_sample (image + dmap), 1539x2051 -> 3x1536x2048", bytes=b, us=t * 1e6, GBps=b / t / 1e9))
    # ground-truth generator: 1000 heads on a 768x1024 image (compute-bound: double exp per pixel and head in range)
    import numpy as np
    rng = np.random.default_rng(0)
    pts = torch.from_numpy(np.stack([rng.random(1000) * 1024, rng.random(1000) * 768], 1)).to(dev)
    t = timeit(lambda: ops.density_map(pts, 768, 1024), iters=5)
    b = 768 * 1024 * 4
    out.append(dict(kernel="density_map, 1000 heads, 768x1024 (FP64-bound, bytes = output only)", bytes=b, us=t * 1e6, GBps=b / t / 1e9))
    for o in out:
        o["frac_of_measured_hbm"] = o["GBps"] / peak
        print("HBM " + json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in o.items()}), flush=True)


if __name__ == "__main__":
    main()


def profile_k1():
    from torch.profiler import ProfilerActivity, profile
    dev = torch.device("cuda:0")
    N, h, w = 16, 96, 128
    cat = ops.new_act(1, N, h, w, 1024, dev, zero=True)
    cat.buf.normal_()
    fv = cat.view(512, 0)
    w1 = torch.randn(4, 512, 512, device=dev) * 0.05
    w2 = torch.randn(4, 512, 512, device=dev) * 0.05
    for _ in range(3):
        ops.ctx_pool(fv, w1, w2)
    torch.cuda.synchronize()
    dense = ops.new_act(1, N, h, w, 512, dev, zero=True)
    dense.buf.normal_()
    for _ in range(2):
        ops.ctx_pool(dense, w1, w2)
    big = ops.new_act(1, N, 2 * h, 2 * w, 512, dev, zero=True)   # 4x the pixels, dense
    for _ in range(2):
        ops.ctx_pool(big, w1, w2)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            ops.ctx_pool(fv, w1, w2)
        torch.cuda.synchronize()
    print("K1PROF --- dense cs=512")
    with profile(activities=[ProfilerActivity.CUDA]) as prof2:
        for _ in range(5):
            ops.ctx_pool(dense, w1, w2)
        torch.cuda.synchronize()
    for e in sorted(prof2.key_averages(), key=lambda e: -e.device_time_total)[:1]:
        print(f"K1PROF dense {e.device_time_total / 5:9.1f} us  {e.key[:50]}")
    with profile(activities=[ProfilerActivity.CUDA]) as prof3:
        for _ in range(5):
            ops.ctx_pool(big, w1, w2)
        torch.cuda.synchronize()
    for e in sorted(prof3.key_averages(), key=lambda e: -e.device_time_total)[:1]:
        print(f"K1PROF big(4x) {e.device_time_total / 5:9.1f} us  {e.key[:50]}")
    for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total):
        if e.device_time_total > 0:
            print(f"K1PROF {e.device_time_total / 5:9.1f} us x{e.count // 5}  {e.key[:70]}")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "k1":
    profile_k1()
