"""GPU parity tests (backward + optimizer step) through the C ABI against the oracle.

Gradient tolerances: a ReLU / max-pool mask is a discontinuity, so a forward error eps flips a
fraction ~eps of the masks and moves the gradient by ~sqrt(eps) in relative L2 (measured: the
fp32 path, forward rel-L2 1.5e-5, gives 1e-5 .. 1.3e-2 per parameter).  BASELINE.json states tolerances for
the forward outputs only.  The bf16 path is held to a yardstick instead of a flat bound: the reference module
ITSELF under CPU bf16 autocast against its own fp64 run, per parameter (tests/golden/make_golden.py --yard:
0.24 on conv1_1's weight falling to 5e-3 at the head) -- the CUDA gradients must stay within 2x of it."""
import os

import numpy as np
import pytest
import torch

from oracle import cannet_torch_cpu as T
from oracle import harness

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _model(sd, precision, dev):
    from can_b200.model import CANNet
    m = CANNet(load_weights=True)
    m.load_state_dict(sd)
    m = m.to(dev)
    m.precision = precision
    return m


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm()).clamp_min(1e-300))


# ---------------------------------------------------------------- single kernels
@pytest.mark.parametrize("case", [(1, 16, 16, 64, 64, 1), (2, 24, 40, 64, 128, 1), (1, 13, 17, 1024, 512, 2),
                                  (1, 24, 32, 128, 64, 2), (1, 16, 8, 256, 256, 1),
                                  (1, 128, 400, 64, 64, 1),    # > 5 patches per CTA: the swapped 64->64 wgrad wraps its stage ring
                                  (2, 13, 21, 64, 64, 1), (1, 11, 19, 128, 128, 1)])   # partial tiles through the staged TMA epilogues
@pytest.mark.parametrize("nsplit", [1, 2])
def test_wgrad_dgrad_bias_match_autograd_oracle(dev, case, nsplit):
    import torch.nn.functional as F
    from can_b200 import ops
    N, H, W, Cin, Cout, dil = case
    g = torch.Generator().manual_seed(11)
    x = torch.relu(torch.randn(N, Cin, H, W, generator=g))
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * (2.0 / (9 * Cin)) ** 0.5
    dz = torch.randn(N, Cout, H, W, generator=g)
    xa, dza = ops.act_from_nchw(x.to(dev), nsplit), ops.act_from_nchw(dz.to(dev), nsplit)
    xr, dzr = xa.to_nchw().cpu().double(), dza.to_nchw().cpu().double()
    wr = w.bfloat16().float()
    if nsplit == 2:
        wr = wr + (w - wr).bfloat16().float()
    xg, wg = xr.clone().requires_grad_(True), wr.double().clone().requires_grad_(True)
    F.conv2d(xg, wg, None, padding=dil, dilation=dil).backward(dzr)   # CPU fp64 truth on the same operands
    grad = torch.empty(Cout, Cin, 3, 3, device=dev)
    ops.conv_wgrad(dza, xa, grad, dil=dil)
    # fp32 accumulate of exact bf16 products; the tensor core's accumulation drifts ~3e-8 per MMA of a chain,
    # so the bound grows with the pixels one split accumulates (51200-pixel case: 1.06e-5 measured)
    assert harness.rel_l2(grad.cpu(), wg.grad) < (1e-5 if N * H * W < 20000 else 2e-5)
    dx = ops.new_act(nsplit, N, H, W, Cin, dev, zero=True)
    ops.conv_tc(dza, ops.pack_conv_weight(w.to(dev), nsplit, dgrad=True), Cin, dil=dil, mode=ops.MODE_STORE,
                relu=False, out=dx, mask=xa)
    ref_dx = xg.grad * (xr > 0)
    assert harness.rel_l2(dx.to_nchw().cpu(), ref_dx) < (3e-3 if nsplit == 1 else 2e-5)
    db = torch.zeros(Cout, device=dev)
    ops.bias_grad(dza, db)
    assert harness.rel_l2(db.cpu(), dzr.sum((0, 2, 3))) < 1e-5


@pytest.mark.parametrize("shape", [(1, 12, 20, 24, 40, 64, 128), (2, 6, 9, 13, 19, 64, 128), (1, 10, 11, 21, 23, 128, 256)])
def test_unpool_routes_gradient_to_first_max(dev, shape):
    """conv (dgrad) + max-unpool epilogue against autograd of max_pool2d(relu(.)) on the CPU; the odd full-resolution
    sizes leave a last row / column that no window covers (must stay zero) and partial tiles for the strided TMA stores."""
    import torch.nn.functional as F
    from can_b200 import ops
    N, H, W, FH, FW, C, C2 = shape   # pooled grid H x W of a FH x FW map (floor-mode 2x2 pooling)
    g = torch.Generator().manual_seed(13)
    pre = torch.randn(N, C, FH, FW, generator=g)
    w_prev = torch.randn(C, C, 3, 3, generator=g) * 0.05
    w = torch.randn(C2, C, 3, 3, generator=g) * 0.05
    dz = torch.randn(N, C2, H, W, generator=g)
    pa = ops.act_from_nchw(pre.to(dev), 1)
    pooled = ops.new_act(1, N, H, W, C, dev)
    arg = torch.empty(N, H, W, C, device=dev, dtype=torch.uint8)
    zero_bias = torch.zeros(C, device=dev)
    ops.conv_tc(pa, ops.pack_conv_weight(w_prev.to(dev), 1), C, mode=ops.MODE_POOL, bias=zero_bias, out=pooled, arg=arg)
    full = ops.new_act(1, N, FH, FW, C, dev)
    ops.conv_tc(pa, ops.pack_conv_weight(w_prev.to(dev), 1), C, mode=ops.MODE_STORE, relu=True, bias=zero_bias, out=full)
    dza = ops.act_from_nchw(dz.to(dev), 1)
    out = ops.new_act(1, N, FH, FW, C, dev, zero=True)
    ops.conv_tc(dza, ops.pack_conv_weight(w.to(dev), 1, dgrad=True), C, mode=ops.MODE_UNPOOL, relu=False, out=out,
                mask=pooled, arg=arg, OH=FH, OW=FW)
    # CPU truth: y = relu(conv_prev) (taken from the device, so ties are identical), X = maxpool(y)
    y = full.to_nchw().cpu().double().requires_grad_(True)
    X = F.max_pool2d(y, 2, 2)
    conv = F.conv2d(X, w.bfloat16().double(), None, padding=1)
    conv.backward(dza.to_nchw().cpu().double())
    ref = y.grad * (y.detach() > 0)
    assert harness.rel_l2(out.to_nchw().cpu(), ref) < 3e-3


# ---------------------------------------------------------------- whole backward
CASES = [("he", 1, 64, 96), ("he", 2, 48, 80), ("default", 1, 104, 136)]
GRAD_TOL = {"fp32": dict(worst=5e-2, median=1e-2, cos=0.999), "bf16": dict(worst=0.35, median=0.12, cos=0.93)}
YARD_FACTOR, YARD_FLOOR = 2.0, 1.5e-2


@pytest.fixture(scope="module")
def yard():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "cannet_bf16_yardstick.npz"))


def test_single_halo_tile_variant_backward(dev, monkeypatch):
    """CAN_B200_HALO1=2: every small-Cin layer (max-unpool epilogues excepted) reads all nine taps out of one halo tile (8-pixel-wide patches, MMA
    descriptors that start off the 1024-byte swizzle pattern) -- same data gradients and parameter gradients."""
    monkeypatch.setenv("CAN_B200_HALO1", "2")
    for case in [(1, 16, 16, 64, 64, 1), (2, 24, 40, 64, 128, 1), (1, 24, 32, 128, 64, 2)]:
        test_wgrad_dgrad_bias_match_autograd_oracle(dev, case, 1)
    test_parameter_gradients_match_oracle(dev, CASES[0], "bf16")


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("case", CASES)
def test_parameter_gradients_match_oracle(dev, case, precision, yard=None):
    regime, n, h, w = case
    if yard is None:
        yard = np.load(os.path.join(os.path.dirname(__file__), "golden", "cannet_bf16_yardstick.npz"))
    sd = harness.make_state_dict(regime)
    x, gt = harness.make_input(n, h, w), harness.make_gt(n, h, w)
    loss_ref, _, gref = T.loss_and_grads(x, gt, sd, dtype=torch.float64)
    m = _model(sd, precision, dev)
    out = m(x.to(dev))
    loss = torch.nn.functional.mse_loss(out, gt.to(dev), reduction="sum")   # stock criterion (train_eval_utils.py:20)
    loss.backward()
    assert float(loss.detach()) == pytest.approx(float(loss_ref), rel=(1e-4 if precision == "fp32" else 2e-2))
    tol = GRAD_TOL[precision]
    rels = {}
    for k, p in m.named_parameters():
        assert p.grad is not None and p.grad.shape == p.shape and torch.isfinite(p.grad).all(), k
        rels[k] = harness.rel_l2(p.grad.cpu(), gref[k])
        assert _cos(p.grad.cpu(), gref[k]) > tol["cos"], (k, rels[k])
        if precision == "bf16":
            # per parameter: within 2x of what bf16 arithmetic does to the reference's own gradient
            y = float(yard[f"{regime}_{n}x{h}x{w}/grad_rel_l2/{k}"])
            assert rels[k] < max(YARD_FACTOR * y, YARD_FLOOR), (k, rels[k], y)
    assert max(rels.values()) < tol["worst"], max(rels.items(), key=lambda kv: kv[1])
    assert float(np.median(list(rels.values()))) < tol["median"]


def test_backward_accumulates_like_autograd(dev):
    """Two backward passes add up in .grad (stock autograd contract used by any torch optimizer)."""
    sd = harness.make_state_dict("he")
    m = _model(sd, "bf16", dev)
    x, gt = harness.make_input(1, 64, 64).to(dev), harness.make_gt(1, 64, 64).to(dev)
    torch.nn.functional.mse_loss(m(x), gt, reduction="sum").backward()
    g1 = m.frontend[21].weight.grad.clone()
    torch.nn.functional.mse_loss(m(x), gt, reduction="sum").backward()
    assert harness.rel_l2(m.frontend[21].weight.grad, 2 * g1) < 1e-5


# ---------------------------------------------------------------- optimizer step
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_fused_train_steps_match_oracle(dev, precision):
    """FusedTrainer (forward, fused loss, backward, fused SGD-momentum on the flat arena) against the
    oracle's restatement of utils/train_eval_utils.py:28-52 + torch.optim.SGD, four steps (two eager, then the
    CUDA-graph capture and a pure replay of the same step)."""
    from can_b200.utils.train_eval_utils import FusedTrainer
    regime, n, h, w = "he", 2, 48, 80
    sd = harness.make_state_dict(regime)
    lr = 1e-7
    ref_sd = {k: v.double().clone() for k, v in sd.items()}
    bufs, ref_losses = {}, []
    m = _model(sd, precision, dev)
    tr = FusedTrainer(m, lr=lr, momentum=0.95)
    losses = []
    for step in range(4):
        x, gt = harness.make_input(n, h, w, seed=100 + step), harness.make_gt(n, h, w, seed=200 + step)
        ref_losses.append(float(T.train_step(x.double(), gt.double(), ref_sd, bufs, lr, 0.95)))
        losses.append(float(tr.step(x.to(dev), gt.to(dev))))
    ltol = 1e-4 if precision == "fp32" else 2e-2
    assert losses == pytest.approx(ref_losses, rel=ltol)
    # parameters moved by the same update (compare the update, not the parameter: lr is tiny)
    cur = m.state_dict()
    utol = 5e-2 if precision == "fp32" else 0.35
    for k in ("backend.10.weight", "backend.0.weight", "conv3_2.weight", "frontend.21.weight", "frontend.7.weight",
              "output_layer.bias"):
        d_ref = (ref_sd[k] - sd[k].double())
        d_got = (cur[k].cpu().double() - sd[k].double())
        assert harness.rel_l2(d_got, d_ref) < utol, k
    assert list(cur.keys()) == list(sd.keys())      # state_dict layout unchanged by the arena


def test_twenty_step_loss_trajectory_follows_the_fp32_reference(dev, yard):
    """20 SGD steps (lr 1e-7, momentum 0.95, a new batch every step) of the CUDA bf16 path against the live reference's
    committed fp32 trajectory (tests/golden/make_golden.py --yard), step by step.  Yardstick: the reference under CPU
    bf16 autocast drifts by up to 1.3 % over the same steps; the CUDA path must stay within 3 % at every step and within
    2x the yardstick on average.  The first two steps are eager, the third is captured, the rest are graph replays."""
    from can_b200.utils.train_eval_utils import FusedTrainer
    ref, ref_bf16 = yard["traj/f32"], yard["traj/bf16"]
    m = _model(harness.make_state_dict("he"), "bf16", dev)
    tr = FusedTrainer(m, lr=1e-7, momentum=0.95)
    got = []
    for step in range(len(ref)):
        x, gt = harness.make_input(2, 48, 80, seed=100 + step), harness.make_gt(2, 48, 80, seed=200 + step)
        got.append(float(tr.step(x.to(dev), gt.to(dev))))
    rel = np.abs(np.array(got) - ref) / ref
    rel_y = np.abs(ref_bf16 - ref) / ref
    print("trajectory rel diff: cuda bf16", np.round(rel, 4), "cpu bf16 yardstick", np.round(rel_y, 4))
    assert rel.max() < 3e-2, (rel, got)
    assert rel.mean() < 2.0 * rel_y.mean() + 2e-3, (rel.mean(), rel_y.mean())


def test_stock_sgd_and_fused_sgd_agree(dev):
    """The module also trains under a stock torch.optim.SGD (reference path, train.py:125-126)."""
    from can_b200.utils.train_eval_utils import FusedTrainer
    sd = harness.make_state_dict("he")
    x, gt = harness.make_input(1, 64, 64).to(dev), harness.make_gt(1, 64, 64).to(dev)
    m1 = _model(sd, "fp32", dev)
    opt = torch.optim.SGD(m1.parameters(), lr=1e-7, momentum=0.95, weight_decay=0)
    m2 = _model(sd, "fp32", dev)
    tr = FusedTrainer(m2, lr=1e-7, momentum=0.95)
    for _ in range(2):
        opt.zero_grad()
        torch.nn.functional.mse_loss(m1(x), gt, reduction="sum").backward()
        opt.step()
        tr.step(x, gt)
    for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
        da, db = a.double().cpu() - sd[k].double(), b.double().cpu() - sd[k].double()
        assert harness.rel_l2(db, da) < 2e-2, k


# ---------------------------------------------------------------- step drivers (reference signatures)
class _Loader(list):
    """Minimal stand-in for the DataLoader the reference passes to train_one_epoch / evaluate."""


def test_train_one_epoch_and_evaluate_mirror_the_reference_drivers(dev):
    from can_b200 import ops
    from can_b200.utils.train_eval_utils import FusedTrainer, evaluate, train_one_epoch
    sd = harness.make_state_dict("he")
    batches = _Loader((harness.make_input(1, 64, 64, seed=10 + i), harness.make_gt(1, 64, 64, seed=20 + i)) for i in range(3))
    # oracle: three SGD steps, running mean of the losses (train_eval_utils.py:41)
    ref_sd, bufs, mean = {k: v.double().clone() for k, v in sd.items()}, {}, 0.0
    for step, (x, gt) in enumerate(batches):
        mean = (mean * step + float(T.train_step(x.double(), gt.double(), ref_sd, bufs, 1e-7, 0.95))) / (step + 1)
    m = _model(sd, "fp32", dev)
    got = train_one_epoch(m, FusedTrainer(m, lr=1e-7, momentum=0.95), batches, dev, epoch=0)
    assert got == pytest.approx(mean, rel=2e-4)
    # evaluate: sum over the loader of |sum(et) - sum(gt)| (train_eval_utils.py:83,136)
    ref_mae = 0.0
    with torch.no_grad():
        for x, gt in batches:
            et = T.forward(x.double(), ref_sd)
            ref_mae += abs(float(et.sum()) - float(gt.double().sum()))
    mae = evaluate(m, batches, dev, epoch=0)
    assert mae == pytest.approx(ref_mae, rel=2e-3)


# ---------------------------------------------------------------- fused first layer (no im2col buffer)
@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 16, 16), (2, 37, 200), (1, 9, 333)])
def test_first_layer_fused_kernels_match_im2col_path(dev, shape):
    """conv1_1 forward and weight gradient rebuilt from the image in shared memory == the im2col + GEMM route
    (same bf16 operands, fp32 accumulation), and both == an fp64 conv of the bf16-rounded operands."""
    import torch.nn.functional as F
    from can_b200 import ops
    N, H, W = shape
    g = torch.Generator().manual_seed(5)
    x = torch.rand(N, 3, H, W, generator=g).to(dev)
    w = (torch.randn(64, 3, 3, 3, generator=g) * 0.2).to(dev)
    b = (torch.randn(64, generator=g) * 0.1).to(dev)
    wp = ops.pack_first_weight(w)
    fused = ops.new_act(1, N, H, W, 64, dev, zero=True)
    ops.conv_first_tc(x, wp, b, fused)
    col = ops.im2col_first(x, 1)
    ref = ops.new_act(1, N, H, W, 64, dev)
    ops.conv_tc(col, wp, 64, taps=1, mode=ops.MODE_STORE, relu=True, bias=b, out=ref)
    assert torch.equal(fused.to_nchw(), ref.to_nchw())
    truth = torch.relu(F.conv2d(x.bfloat16().double().cpu(), w.bfloat16().double().cpu(), b.double().cpu(), padding=1))
    assert harness.rel_l2(fused.to_nchw().cpu(), truth) < 3e-3          # bf16 output rounding
    # weight gradient
    dz = torch.randn(N, 64, H, W, generator=g)
    dza = ops.act_from_nchw(dz.to(dev), 1)
    g_fused = torch.empty(64, 3, 3, 3, device=dev)
    ops.conv_first_wgrad(dza, x, g_fused)
    xg = x.bfloat16().double().cpu()
    wg = torch.zeros(64, 3, 3, 3, dtype=torch.float64, requires_grad=True)
    F.conv2d(xg, wg, None, padding=1).backward(dza.to_nchw().cpu().double())
    assert harness.rel_l2(g_fused.cpu(), wg.grad) < 1e-5
    g_acc = g_fused.clone()
    ops.conv_first_wgrad(dza, x, g_acc, accumulate=True, alpha=0.5)
    assert harness.rel_l2(g_acc.cpu(), 1.5 * wg.grad) < 1e-5


@pytest.mark.parametrize("hw", [(13, 17), (48, 64), (9, 21), (96, 128)])
def test_pool_transpose_operands_match_oracle(dev, hw):
    """can_ctx_pint / can_ctx_dtab (the A and B operands of MODE_DFV's extra k-block): bit-exact against the oracle's
    pooling-transpose matrix rounded to bf16, and a plain transpose + rounding of the pooled-map gradient."""
    from can_b200 import ops
    from oracle import cannet_oracle as O
    h, w = hw
    got = ops.ctx_pint(h, w, dev).float().cpu()
    want = torch.zeros(h, w, 64)
    want[:, :, :50] = -torch.from_numpy(O.pool_transpose_matrix(h, w, dtype=np.float32))
    assert torch.equal(got, want.bfloat16().float())
    g = torch.Generator().manual_seed(6)
    dpool = torch.randn(3, 50, 512, generator=g)
    d = ops.ctx_dtab(dpool.to(dev)).float().cpu()
    ref = torch.zeros(3, 512, 64)
    ref[:, :, :50] = dpool.transpose(1, 2)
    assert torch.equal(d, ref.bfloat16().float())


def test_transpose_multi_is_exact(dev):
    """can_transpose_f32_multi (the eight W^T tables of the contextual backward in one launch) against torch."""
    from can_b200 import ops
    g = torch.Generator().manual_seed(7)
    ws = [torch.randn(512, 512, 1, 1, generator=g).to(dev) for _ in range(8)]
    got = ops.transpose_stack(ws)
    torch.cuda.synchronize()
    assert got.shape == (8, 512, 512)
    for j, w in enumerate(ws):
        assert torch.equal(got[j], w.reshape(512, 512).t())
