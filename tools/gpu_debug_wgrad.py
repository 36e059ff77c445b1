"""GPU bring-up probe: wgrad / dgrad / unpool kernels in isolation against torch autograd (fp64)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))

CASES = [
    # name, N, H, W, Cin, Cout, dil, nsplit
    ("w64_64", 1, 16, 16, 64, 64, 1, 1),
    ("w64_128", 2, 24, 40, 64, 128, 1, 1),
    ("w256_512", 1, 16, 32, 256, 512, 1, 1),
    ("w1024_512_d2", 1, 13, 17, 1024, 512, 2, 1),
    ("w128_64_d2", 1, 24, 32, 128, 64, 2, 1),
    ("w128_128_s3", 1, 16, 16, 128, 128, 1, 3),
    ("w512_256_s3", 1, 24, 32, 512, 256, 2, 3),
]


def run_case(name):
    import torch
    import torch.nn.functional as F
    from can_b200 import ops
    _, N, H, W, Cin, Cout, dil, nsplit = [c for c in CASES if c[0] == name][0]
    g = torch.Generator().manual_seed(11)
    x = torch.relu(torch.randn(N, Cin, H, W, generator=g)).cuda()
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) * (2.0 / (9 * Cin)) ** 0.5).cuda()
    dz = torch.randn(N, Cout, H, W, generator=g).cuda()
    xa, dza = ops.act_from_nchw(x, nsplit), ops.act_from_nchw(dz, nsplit)
    xr, dzr = xa.to_nchw().double(), dza.to_nchw().double()
    wr = w.bfloat16().float()
    if nsplit >= 2:
        r1 = w - wr
        wr = wr + r1.bfloat16().float()
        if nsplit == 3:
            wr = wr + (w - wr).bfloat16().float()
    wr = wr.double()
    xg = xr.clone().requires_grad_(True)
    wg = wr.clone().requires_grad_(True)
    y = F.conv2d(xg, wg, None, padding=dil, dilation=dil)
    y.backward(dzr)
    res = {"name": name}
    grad = torch.empty(Cout, Cin, 3, 3, device="cuda")
    ops.conv_wgrad(dza, xa, grad, dil=dil)
    res["wgrad_rel"] = float((grad.double() - wg.grad).norm() / wg.grad.norm())
    wd = ops.pack_conv_weight(w, nsplit, dgrad=True)
    dx = ops.new_act(nsplit, N, H, W, Cin, x.device, zero=True)
    ops.conv_tc(dza, wd, Cin, dil=dil, mode=ops.MODE_STORE, relu=False, out=dx, mask=xa)
    ref_dx = xg.grad * (xr > 0)
    res["dgrad_rel"] = float((dx.to_nchw().double() - ref_dx).norm() / ref_dx.norm())
    db = torch.zeros(Cout, device="cuda")
    ops.bias_grad(dza, db)
    res["bias_rel"] = float((db.double() - dzr.sum((0, 2, 3))).norm() / dzr.sum((0, 2, 3)).norm())
    torch.cuda.synchronize()
    if res["wgrad_rel"] > 2e-2:
        e = (grad.double() - wg.grad).abs()
        res["err_by_tap"] = [round(float(v), 4) for v in e.mean(dim=(0, 1)).flatten()]
        res["err_by_co16"] = [round(float(v), 4) for v in e.mean(dim=(1, 2, 3))[:64].view(-1, 16).mean(1)]
        res["err_by_ci16"] = [round(float(v), 4) for v in e.mean(dim=(0, 2, 3))[:64].view(-1, 16).mean(1)]
        res["got_norm"], res["ref_norm"] = float(grad.norm()), float(wg.grad.norm())
    print("RESULT " + json.dumps(res), flush=True)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--case":
        run_case(sys.argv[2])
        return
    for n in (sys.argv[1:] or [c[0] for c in CASES]):
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", n], capture_output=True, text=True,
                               timeout=180)
            lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            print(n, "rc=%d" % r.returncode, lines[-1] if lines else ("NO RESULT; stderr tail: " + r.stderr[-800:]),
                  flush=True)
        except subprocess.TimeoutExpired:
            print(n, "TIMEOUT", flush=True)


if __name__ == "__main__":
    main()
