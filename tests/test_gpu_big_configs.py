"""GPU parity at BASELINE.json's own shapes (configs[1..4]): 3x512x512, 3x768x1024, 3x2048x2048, batches 4 / 8 / 16.

The small-shape tests run most persistent loops once; here every CTA walks many tiles, K1 streams more rows than
there are CTAs, split-K factors are the ones the cost model picks for the benchmarked step, and the pooling windows
overlap on both axes (256 is not divisible by 3 or 6).  Truth: the committed outputs of the live reference module
(``tests/golden/cannet_golden_big.npz``, made by ``make_golden.py --big``) and, for batches and training steps, the
oracle's torch-CPU port run in the test on this box's host cores.
"""
import os

import numpy as np
import pytest
import torch

from oracle import cannet_oracle as O
from oracle import cannet_torch_cpu as T
from oracle import harness

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# BASELINE.json north_star: fp32 density rel-L2 1e-4 / count 1e-3; bf16 count 1e-2 (bf16 rel-L2: ~2x the reference's own
# CPU-bf16 error)
TOL = {"fp32": dict(l2=1e-4, count=1e-3), "bf16": dict(l2=3e-2, count=1e-2)}
BIG = {"he_1x512x512": (512, 512), "he_1x768x1024": (768, 1024), "he_1x2048x2048": (2048, 2048)}


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def big():
    return np.load(os.path.join(ROOT, "tests", "golden", "cannet_golden_big.npz"))


@pytest.fixture(scope="module")
def sd():
    return harness.make_state_dict("he")


def _model(sd, precision, dev):
    from can_b200.model import CANNet
    m = CANNet(load_weights=True)
    m.load_state_dict(sd)
    m = m.to(dev)
    m.precision = precision
    return m


# ---------------------------------------------------------------- forward, one image, all three sizes
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(BIG))
def test_forward_matches_reference_at_baseline_sizes(dev, big, sd, name, precision):
    h, w = BIG[name]
    m = _model(sd, precision, dev)
    with torch.no_grad():
        out = m(harness.make_input(1, h, w).to(dev)).cpu()
    ref = torch.from_numpy(big[f"{name}/f64/out"])
    assert out.shape == ref.shape
    assert harness.rel_l2(out, ref) < TOL[precision]["l2"]
    assert harness.rel_count(out, ref) < TOL[precision]["count"]


# ---------------------------------------------------------------- inference batch 16 (configs[1]) and batch 8 at 2048 (configs[4])
@pytest.mark.parametrize("cfg", [("bf16", 16, 768, 1024), ("fp32", 16, 768, 1024), ("bf16", 4, 2048, 2048)])
def test_batched_inference_counts_match_single_image_oracle(dev, sd, cfg):
    """Every image of a batch against the port run on it alone: count within the north_star tolerance, map within the
    rel-L2 bound (16 distinct images; the 2048 case checks the first and last image only -- 40 s of CPU each otherwise)."""
    from can_b200 import ops
    precision, n, h, w = cfg
    x = harness.make_input(n, h, w, seed=1234)
    gt = harness.make_gt(n, h, w)
    m = _model(sd, precision, dev)
    with torch.no_grad():
        out = m(x.to(dev))
        _, ce, cg, _ = ops.mse_count(out, gt.to(dev), want_grad=False)
    out, ce, cg = out.cpu(), ce.cpu(), cg.cpu()
    torch.set_num_threads(os.cpu_count() or 1)
    which = range(n) if h * w <= 768 * 1024 else (0, n - 1)
    with torch.no_grad():
        for i in which:
            ref = T.forward(x[i:i + 1], sd)
            assert harness.rel_l2(out[i:i + 1], ref) < max(TOL[precision]["l2"], 2e-5), i   # (fp32 port: ~1e-6 of its own noise)
            assert abs(float(ce[i]) - float(ref.double().sum())) <= TOL[precision]["count"] * abs(float(ref.double().sum())), i
            assert float(cg[i]) == pytest.approx(float(gt[i].double().sum()), rel=1e-5)


# ---------------------------------------------------------------- K1 with more rows than CTAs, overlapping windows
@pytest.mark.parametrize("shape", [(4, 96, 128), (16, 96, 128), (2, 256, 256), (3, 64, 64)])
def test_ctx_pool_rows_loop_and_ring_wrap(dev, shape):
    """N*H = 384 / 1536 / 512 rows on 148 persistent CTAs: the per-CTA row loop and the TMA ring wrap several times;
    256 and 64 are not divisible by 3 or 6 (windows overlap by one row / column)."""
    from can_b200 import ops
    n, h, w = shape
    g = torch.Generator().manual_seed(21)
    fv = torch.rand(n, 512, h, w, generator=g)
    w1 = [torch.randn(512, 512, 1, 1, generator=g) * 0.05 for _ in range(4)]
    w2 = [torch.randn(512, 512, 1, 1, generator=g) * 0.05 for _ in range(4)]
    fa = ops.act_from_nchw(fv.to(dev), 1)
    fr = fa.to_nchw().cpu().double()
    pooled, tabA, tabB = ops.ctx_pool(fa, ops.stack_1x1([t.to(dev) for t in w1]), ops.stack_1x1([t.to(dev) for t in w2]))
    off = 0
    for j, s in enumerate(O.SCALES):
        ref = torch.nn.functional.adaptive_avg_pool2d(fr, (s, s))              # ATen, fp64 (== O.adaptive_avg_pool, CPU tests)
        if n * h * w <= 4 * 96 * 128:
            assert np.abs(O.adaptive_avg_pool(fr.numpy(), s) - ref.numpy()).max() < 1e-12
        got = pooled[:, off:off + s * s].cpu().double()                         # [N, s*s, 512]
        assert (got - ref.reshape(n, 512, s * s).transpose(1, 2)).abs().max() < 1e-5
        a_ref = torch.nn.functional.conv2d(ref, w1[j].double())
        b_ref = torch.nn.functional.conv2d(a_ref, w2[j].double())
        assert (tabA[:, off:off + s * s].cpu().double() - a_ref.reshape(n, 512, s * s).transpose(1, 2)).abs().max() < 1e-4
        assert (tabB[:, off:off + s * s].cpu().double() - b_ref.reshape(n, 512, s * s).transpose(1, 2)).abs().max() < 1e-4
        off += s * s


# ---------------------------------------------------------------- contextual module at 96x128 / 256x256 against the port
@pytest.mark.parametrize("cfg", [("bf16", 2, 768, 1024, 3e-2), ("fp32", 1, 768, 1024, 5e-5), ("bf16", 1, 2048, 2048, 3e-2)])
def test_fv_and_fi_match_port_at_baseline_sizes(dev, sd, cfg):
    """fv (conv4_3 output) and fi (MODE_CTX2 / MODE_CTX output) of the engine against the port's intermediates."""
    from can_b200 import engine
    precision, n, h, w, tol = cfg
    x = harness.make_input(n, h, w, seed=99)
    m = _model(sd, precision, dev)
    with torch.no_grad():
        _, saved = engine.forward(m, x.to(dev), m.nsplit, save=True)
    fv = saved["cat"].view(512, 0).to_nchw().cpu()
    fi = saved["cat"].view(512, 512).to_nchw().cpu()
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.no_grad():
        _, inter = T.forward(x, sd, return_intermediates=True)
    assert harness.rel_l2(fv, inter["fv"]) < tol
    assert harness.rel_l2(fi, inter["fi"]) < tol


# ---------------------------------------------------------------- gradients at 768x1024 / 512x512 against the live reference's
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["he_1x512x512", "he_1x768x1024"])
def test_parameter_gradient_summaries_match_reference(dev, big, sd, name, precision):
    """Stock autograd route (loss.backward() on the module): per parameter the gradient's sum, L2 norm and first five
    entries against the fp64 reference's (summaries only are committed: 83 MB of gradients per case otherwise)."""
    h, w = BIG[name]
    m = _model(sd, precision, dev)
    x, gt = harness.make_input(1, h, w).to(dev), harness.make_gt(1, h, w).to(dev)
    loss = torch.nn.functional.mse_loss(m(x), gt, reduction="sum")
    loss.backward()
    assert float(loss.detach()) == pytest.approx(float(big[f"{name}/f64/loss"]), rel=1e-4 if precision == "fp32" else 2e-2)
    ntol = 2e-3 if precision == "fp32" else 8e-2
    worst = 0.0
    for k, p in m.named_parameters():
        ref = big[f"{name}/grad/{k}"]     # [sum, norm, absmax, first five]
        g = p.grad.detach().double().flatten().cpu()
        assert torch.isfinite(g).all(), k
        rn = abs(float(g.norm()) - ref[1]) / ref[1]
        worst = max(worst, rn)
        assert rn < ntol, (k, rn)
        # the first five entries, on the scale of the largest entry (single entries of a gradient can be ~0)
        d5 = float((g[:5] - torch.from_numpy(ref[3:8])).abs().max()) / ref[2]
        assert d5 < (2e-3 if precision == "fp32" else 5e-2), (k, d5)
    print(f"{name} {precision}: worst gradient-norm error {worst:.2e}")


# ---------------------------------------------------------------- the benchmarked training steps (configs[2], configs[3])
@pytest.mark.parametrize("cfg", [(4, 768, 1024), (8, 512, 512)])
def test_fused_train_step_matches_port_at_benchmarked_config(dev, sd, cfg):
    """FusedTrainer.step (bf16) on the bench's own batch against the port's restatement of train_eval_utils.py:28-52 +
    SGD run on this box's CPU: loss, and the applied update of one parameter per gradient bucket.  Three steps (two
    eager, then the captured graph), the last compared on loss only."""
    from can_b200.utils.train_eval_utils import FusedTrainer
    n, h, w = cfg
    lr = 1e-10     # (bench.py's rate: He-initialised weights under a sum loss diverge at the reference's 1e-7)
    x, gt = harness.make_input(n, h, w, seed=1234), harness.make_gt(n, h, w, seed=4321)
    torch.set_num_threads(os.cpu_count() or 1)
    ref_sd = {k: v.clone() for k, v in sd.items()}
    bufs = {}
    ref_losses = [float(T.train_step(x, gt, ref_sd, bufs, lr, 0.95))]
    m = _model(sd, "bf16", dev)
    tr = FusedTrainer(m, lr=lr, momentum=0.95)
    xd, gd = x.to(dev), gt.to(dev)
    losses = [float(tr.step(xd, gd))]
    assert losses[0] == pytest.approx(ref_losses[0], rel=2e-2)
    cur = m.state_dict()
    for k in ("backend.10.weight", "backend.0.weight", "output_layer.weight", "conv3_2.weight", "conv6_1.weight",
              "frontend.21.weight", "frontend.14.weight", "frontend.2.weight", "frontend.0.weight", "backend.4.bias"):
        d_ref = ref_sd[k].double() - sd[k].double()
        d_got = cur[k].cpu().double() - sd[k].double()
        # fp32 parameters hold ~7 digits: an update of lr * g is resolved to ~1e-7 |p| / (lr |g|); compare on the update
        # applied to the parameter, as the caller sees it
        assert harness.rel_l2(d_got, d_ref) < 0.35, (k, harness.rel_l2(d_got, d_ref))
        cos = float((d_got.flatten() @ d_ref.flatten()) / (d_got.norm() * d_ref.norm()).clamp_min(1e-300))
        assert cos > 0.93, (k, cos)
    for _ in range(3):     # eager, capture, replay: the loss keeps following the port's trajectory
        ref_losses.append(float(T.train_step(x, gt, ref_sd, bufs, lr, 0.95)))
        losses.append(float(tr.step(xd, gd)))
    # The same batch four times with momentum: the loss falls ~10x over these steps, so a 0.5 % forward difference of
    # the bf16 path is amplified step by step (measured 2.3 % / 2.8 % at steps 3 / 4); the bound widens accordingly.
    for k, (a, b) in enumerate(zip(losses, ref_losses)):
        assert a == pytest.approx(b, rel=2e-2 * (1 + k)), (k, losses, ref_losses)
