"""Generate golden vectors from the LIVE reference module.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python tests/golden/make_golden.py          # small cases  -> cannet_golden.npz
    python tests/golden/make_golden.py --big    # BASELINE.json's own shapes -> cannet_golden_big.npz
    python tests/golden/make_golden.py --yard   # bf16 yardstick of the reference itself -> cannet_bf16_yardstick.npz

The reference is path-loaded (``model/__init__.py`` imports matplotlib, absent here)
and built with ``CANNet(load_weights=True)`` (the default constructor downloads
VGG-16).  Weights come from ``oracle.harness.make_state_dict``
so fixtures stay small: only outputs, checksums and gradient summaries are stored.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import harness  # noqa: E402

CASES = [
    # name, regime, N, H, W
    ("he_1x64x96", "he", 1, 64, 96),
    ("he_2x48x80", "he", 2, 48, 80),
    ("he_1x104x136", "he", 1, 104, 136),
    ("default_1x64x64", "default", 1, 64, 64),
    ("ref_1x64x64", "ref", 1, 64, 64),
    ("he_1x128x192", "he", 1, 128, 192),
]


# BASELINE.json configs at their own image sizes (one image each; batches are checked image by image).
# Stored: the fp64 and fp32 density maps, the loss and the per-parameter gradient summaries -- no features.
BIG_CASES = [
    # name, regime, N, H, W, backward?
    ("he_1x512x512", "he", 1, 512, 512, True),       # configs[2]
    ("he_1x768x1024", "he", 1, 768, 1024, True),     # configs[1], configs[3]
    ("he_1x2048x2048", "he", 1, 2048, 2048, False),  # configs[4] (inference only: forward, fp64)
]


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_cannet", "/root/reference/model/CANNet.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def summarize(t):
    t = t.detach().double().flatten()
    return np.array([t.sum().item(), t.norm().item(), t.abs().max().item()] + t[:5].tolist())


def main_big():
    torch.set_num_threads(os.cpu_count())
    ref = load_reference()
    out = {}
    for name, regime, n, h, w, bwd in BIG_CASES:
        sd = harness.make_state_dict(regime, seed=0)
        x = harness.make_input(n, h, w)
        gt = harness.make_gt(n, h, w)
        for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            if not bwd and tag == "f32":
                continue
            m = ref.CANNet(load_weights=True).to(dt)
            m.load_state_dict({k: v.to(dt) for k, v in sd.items()}, strict=True)
            with torch.set_grad_enabled(bwd):
                et = m(x.to(dt))
                loss = torch.nn.MSELoss(reduction="sum")(et, gt.to(dt))  # train_eval_utils.py:20,37
            if bwd:
                loss.backward()
            out[f"{name}/{tag}/out"] = et.detach().numpy()
            out[f"{name}/{tag}/loss"] = np.array(loss.item())
            if bwd and tag == "f64":
                for k, p in m.state_dict(keep_vars=True).items():
                    out[f"{name}/grad/{k}"] = summarize(p.grad)
            del m, et, loss
        print(name, "ok", out[f"{name}/f64/out"].shape, float(out[f"{name}/f64/loss"]), flush=True)
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "cannet_golden_big.npz"), **out)


YARD_CASES = [("he_1x64x96", "he", 1, 64, 96), ("he_2x48x80", "he", 2, 48, 80), ("default_1x104x136", "default", 1, 104, 136),
              ("he_1x256x256", "he", 1, 256, 256)]
TRAJ = dict(regime="he", n=2, h=48, w=80, steps=20, lr=1e-7, momentum=0.95)


def main_yard():
    """How far the REFERENCE ITSELF moves when its convolutions run in bf16 (torch.autocast on the CPU: bf16 weights and
    activations, fp32 accumulation -- the arithmetic of this repo's bf16 path): per-parameter gradient error against its
    own fp64 run, and the loss trajectory of 20 SGD steps (train_eval_utils.py:28-52, train.py:63,125-126) against fp32.
    The GPU tests hold the CUDA bf16 path to a small multiple of these figures instead of a flat tolerance."""
    torch.set_num_threads(os.cpu_count())
    ref = load_reference()
    out = {}
    crit = torch.nn.MSELoss(reduction="sum")
    for name, regime, n, h, w in YARD_CASES:
        sd = harness.make_state_dict(regime, seed=0)
        x, gt = harness.make_input(n, h, w), harness.make_gt(n, h, w)
        m64 = ref.CANNet(load_weights=True).double()
        m64.load_state_dict({k: v.double() for k, v in sd.items()})
        crit(m64(x.double()), gt.double()).backward()
        m = ref.CANNet(load_weights=True)
        m.load_state_dict(sd)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            et = m(x)
        loss = crit(et.float(), gt)
        loss.backward()
        out[f"{name}/fwd_rel_l2"] = np.array(harness.rel_l2(et.float().detach(), m64(x.double()).detach()))
        for (k, p), (_, p64) in zip(m.named_parameters(), m64.named_parameters()):
            out[f"{name}/grad_rel_l2/{k}"] = np.array(harness.rel_l2(p.grad, p64.grad))
        worst = max(float(v) for kk, v in out.items() if kk.startswith(f"{name}/grad_rel_l2/"))
        print(name, "fwd", float(out[f"{name}/fwd_rel_l2"]), "worst grad", worst, flush=True)
    # loss trajectories: fp32 and bf16-autocast runs of the reference module under torch.optim.SGD
    t = TRAJ
    sd = harness.make_state_dict(t["regime"], seed=0)
    traj = {}
    for tag in ("f32", "bf16"):
        m = ref.CANNet(load_weights=True)
        m.load_state_dict(sd)
        opt = torch.optim.SGD(m.parameters(), lr=t["lr"], momentum=t["momentum"], weight_decay=0)   # train.py:63,125-126
        losses = []
        for step in range(t["steps"]):
            x = harness.make_input(t["n"], t["h"], t["w"], seed=100 + step)
            gt = harness.make_gt(t["n"], t["h"], t["w"], seed=200 + step)
            opt.zero_grad()
            if tag == "bf16":
                with torch.autocast("cpu", dtype=torch.bfloat16):
                    et = m(x)
                et = et.float()
            else:
                et = m(x)
            loss = crit(et, gt)
            loss.backward()
            opt.step()
            losses.append(loss.item())
        traj[tag] = np.array(losses)
    out["traj/f32"], out["traj/bf16"] = traj["f32"], traj["bf16"]
    print("trajectory f32 ", traj["f32"])
    print("trajectory bf16 rel diff", np.abs(traj["bf16"] - traj["f32"]) / traj["f32"])
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "cannet_bf16_yardstick.npz"), **out)


def main():
    torch.set_num_threads(os.cpu_count())
    ref = load_reference()
    out = {}
    for name, regime, n, h, w in CASES:
        sd = harness.make_state_dict(regime, seed=0)
        x = harness.make_input(n, h, w)
        gt = harness.make_gt(n, h, w)
        for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            m = ref.CANNet(load_weights=True).to(dt)
            m.load_state_dict({k: v.to(dt) for k, v in sd.items()}, strict=True)
            feats = {}
            m.frontend.register_forward_hook(lambda mod, i, o: feats.__setitem__("fv", o.detach().clone()))
            m.backend.register_forward_hook(lambda mod, i, o: feats.__setitem__("cat", i[0].detach().clone()))
            et = m(x.to(dt))
            loss = torch.nn.MSELoss(reduction="sum")(et, gt.to(dt))  # train_eval_utils.py:20,37
            loss.backward()
            out[f"{name}/{tag}/out"] = et.detach().numpy()
            out[f"{name}/{tag}/loss"] = np.array(loss.item())
            if tag == "f64":
                out[f"{name}/fv_sum"] = summarize(feats["fv"])
                out[f"{name}/fi_sum"] = summarize(feats["cat"][:, 512:])
                # full fv / fi only for the smallest cases (kept < 1 MB each)
                if h * w <= 64 * 96:
                    out[f"{name}/fv"] = feats["fv"].float().numpy()
                    out[f"{name}/fi"] = feats["cat"][:, 512:].float().numpy()
                for k, p in m.state_dict(keep_vars=True).items():
                    out[f"{name}/grad/{k}"] = summarize(p.grad)
        print(name, "ok", out[f"{name}/f64/out"].shape, float(out[f"{name}/f64/loss"]))
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "cannet_golden.npz"), **out)


if __name__ == "__main__":
    main_big() if "--big" in sys.argv[1:] else main_yard() if "--yard" in sys.argv[1:] else main()
