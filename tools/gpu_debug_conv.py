"""GPU bring-up probe for the tcgen05 implicit-GEMM kernel (not a pytest; run under gpurun).

Each case runs in its own subprocess with a timeout so a trap or hang in one case cannot
take the others (or the box) down.  The checker is torch's fp64 conv on the same device --
test infrastructure only.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))

CASES = [
    # name, N, H, W, Cin, Cout, dil, mode, bn, nsplit
    ("c64_64_s", 1, 16, 16, 64, 64, 1, "store", 0, 1),
    ("c64_128_s", 1, 16, 32, 64, 128, 1, "store", 0, 1),
    ("c128_256_s", 2, 24, 40, 128, 256, 1, "store", 0, 1),
    ("c512_512_bn128", 1, 16, 16, 512, 512, 1, "store", 128, 1),
    ("c512_512_bn256", 1, 16, 16, 512, 512, 1, "store", 256, 1),
    ("c1024_512_d2", 1, 24, 32, 1024, 512, 2, "store", 0, 1),
    ("c64_64_pool", 2, 32, 48, 64, 64, 1, "pool", 0, 1),
    ("c256_256_pool_odd", 1, 26, 34, 256, 256, 1, "pool", 0, 1),
    ("c128_64_head", 1, 13, 17, 128, 64, 2, "head", 0, 1),
    ("c64_64_w8", 1, 16, 8, 64, 64, 1, "store", 0, 1),
    ("c128_128_split", 1, 16, 16, 128, 128, 1, "store", 0, 2),
    ("c512_512_split", 1, 24, 32, 512, 512, 2, "store", 0, 2),
]


def run_case(name):
    import torch
    import torch.nn.functional as F
    from can_b200 import ops
    case = [c for c in CASES if c[0] == name][0]
    _, N, H, W, Cin, Cout, dil, mode, bn, nsplit = case
    g = torch.Generator().manual_seed(7)
    x = torch.randn(N, Cin, H, W, generator=g).cuda()
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) * (2.0 / (9 * Cin)) ** 0.5).cuda()
    b = (torch.randn(Cout, generator=g) * 0.1).cuda()
    xa = ops.act_from_nchw(x, nsplit)
    xr = xa.to_nchw()  # the value the kernel really sees (bf16 or split)
    wp = ops.pack_conv_weight(w, nsplit)
    if nsplit == 1:
        wr = w.bfloat16().float()
    else:
        hi = w.bfloat16().float()
        wr = hi + (w - hi).bfloat16().float()
    ref = F.relu(F.conv2d(xr.double(), wr.double(), b.double(), padding=dil, dilation=dil))
    res = {"name": name}
    if mode == "store":
        out = ops.new_act(nsplit, N, H, W, Cout, x.device, zero=True)
        ops.conv_tc(xa, wp, Cout, dil=dil, mode=ops.MODE_STORE, relu=True, bias=b, out=out, bn=bn)
        got = out.to_nchw().double()
    elif mode == "pool":
        out = ops.new_act(nsplit, N, H // 2, W // 2, Cout, x.device, zero=True)
        arg = torch.full((N, H // 2, W // 2, Cout), 9, device=x.device, dtype=torch.uint8)
        ops.conv_tc(xa, wp, Cout, dil=dil, mode=ops.MODE_POOL, bias=b, out=out, arg=arg, bn=bn)
        got = out.to_nchw().double()
        refb = ref.float().bfloat16().double() if nsplit == 1 else ref
        pooled, idx = F.max_pool2d(refb, 2, 2, return_indices=True)
        ref = pooled
        code = arg.permute(0, 3, 1, 2).long()
        yy = torch.arange(H // 2, device=x.device).view(1, 1, -1, 1) * 2 + code // 2
        xx = torch.arange(W // 2, device=x.device).view(1, 1, 1, -1) * 2 + code % 2
        res["argmax_mismatch"] = int(((yy * W + xx) != idx).sum().item())
        res["argmax_total"] = idx.numel()
        res["argmax_maxcode"] = int(arg.max().item())
    elif mode == "head":
        out = ops.new_act(nsplit, N, H, W, Cout, x.device, zero=True)
        hw = (torch.randn(64, generator=g) * 0.2).cuda()
        hb = torch.tensor([0.3]).cuda()
        dens = torch.zeros(N, 1, H, W, device=x.device)
        ops.conv_tc(xa, wp, Cout, dil=dil, mode=ops.MODE_HEAD, bias=b, out=out, head_w=hw, head_b=hb, head_out=dens)
        got = out.to_nchw().double()
        dref = (ref * hw.double().view(1, 64, 1, 1)).sum(1, keepdim=True) + 0.3
        res["head_rel"] = float((dens.double() - dref).norm() / dref.norm())
    torch.cuda.synchronize()
    err = (got - ref)
    res["rel_l2"] = float(err.norm() / ref.norm())
    res["max_abs"] = float(err.abs().max())
    res["ref_norm"] = float(ref.norm())
    res["got_norm"] = float(got.norm())
    if res["rel_l2"] > 2e-2:
        # where is it wrong?  per-channel-block / per-row / per-column error maps decode layout bugs
        e = err.abs()
        cb = min(64, e.shape[1])
        res["err_by_cblock"] = [round(float(v), 4) for v in e.mean(dim=(0, 2, 3)).view(-1, cb).mean(1)[:8]]
        res["err_by_ch16"] = [round(float(v), 4) for v in e.mean(dim=(0, 2, 3))[:64].view(-1, 16).mean(1)]
        res["err_by_row"] = [round(float(v), 4) for v in e.mean(dim=(0, 1, 3))[:16]]
        res["err_by_col"] = [round(float(v), 4) for v in e.mean(dim=(0, 1, 2))[:16]]
        res["sample_got"] = [round(float(v), 4) for v in got[0, :4, 0, :4].flatten()]
        res["sample_ref"] = [round(float(v), 4) for v in ref[0, :4, 0, :4].flatten()]
    print("RESULT " + json.dumps(res), flush=True)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--case":
        run_case(sys.argv[2])
        return
    names = sys.argv[1:] or [c[0] for c in CASES]
    for n in names:
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", n], capture_output=True, text=True,
                               timeout=180)
            lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            print(n, "rc=%d" % r.returncode, lines[-1] if lines else ("NO RESULT; stderr tail: " + r.stderr[-600:]),
                  flush=True)
        except subprocess.TimeoutExpired:
            print(n, "TIMEOUT", flush=True)


if __name__ == "__main__":
    main()
