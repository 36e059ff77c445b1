"""GPU parity tests (forward path): every check goes through the C ABI of libcan_b200.so and
compares with the oracle (numpy / torch-CPU restatement) or the committed golden vectors."""
import numpy as np
import pytest
import torch

from oracle import cannet_oracle as O
from oracle import harness

pytestmark = pytest.mark.gpu

CASES = {
    "he_1x64x96": ("he", 1, 64, 96),
    "he_2x48x80": ("he", 2, 48, 80),
    "he_1x104x136": ("he", 1, 104, 136),
    "default_1x64x64": ("default", 1, 64, 64),
    "ref_1x64x64": ("ref", 1, 64, 64),
    "he_1x128x192": ("he", 1, 128, 192),
}
# tolerances from BASELINE.json north_star: fp32 density rel-L2 1e-4 / count 1e-3; bf16 count 1e-2.
# (bf16 density rel-L2 is not specified there; 3e-2 is ~2x the reference's own CPU-bf16 error)
TOL = {"fp32": dict(l2=1e-4, count=1e-3), "bf16": dict(l2=3e-2, count=1e-2)}


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _model(regime, precision, dev):
    from can_b200.model import CANNet
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict(regime))
    m = m.to(dev)
    m.precision = precision
    return m


# ---------------------------------------------------------------- integer tables: bit-exact
@pytest.mark.parametrize("in_size", [1, 2, 3, 6])
def test_bilinear_taps_bit_exact(dev, in_size):
    from can_b200 import ops
    for out_size in [1, 2, 5, 6, 7, 8, 13, 17, 48, 64, 96, 128, 255, 256]:
        i0, i1, l0, l1 = [t.cpu().numpy() for t in ops.bilinear_taps(in_size, out_size, dev)]
        r0, r1, rl0, rl1 = O.bilinear_taps(in_size, out_size)
        assert np.array_equal(i0, r0) and np.array_equal(i1, r1), (in_size, out_size)
        assert np.array_equal(l0.view(np.uint32), rl0.view(np.uint32)), (in_size, out_size)
        assert np.array_equal(l1.view(np.uint32), rl1.view(np.uint32)), (in_size, out_size)


@pytest.mark.parametrize("s", [1, 2, 3, 6])
def test_adaptive_windows_bit_exact(dev, s):
    from can_b200 import ops
    for in_size in [1, 5, 6, 7, 8, 13, 17, 48, 64, 96, 128, 256]:
        st, en = [t.cpu().numpy().tolist() for t in ops.adaptive_windows(in_size, s, dev)]
        assert list(zip(st, en)) == O.adaptive_windows(in_size, s)


# ---------------------------------------------------------------- tcgen05 conv vs the numpy oracle
def _bf16(x):
    return x.bfloat16().float()


CONV_CASES = [
    # N, H, W, Cin, Cout, dil, bn
    (1, 16, 16, 64, 64, 1, 0),
    (2, 24, 40, 128, 256, 1, 0),
    (1, 16, 16, 512, 512, 1, 128),
    (1, 24, 32, 1024, 512, 2, 256),
    (1, 13, 17, 256, 128, 2, 0),   # ragged tile (H, W not multiples of the 8x16 patch)
    (1, 16, 8, 64, 64, 1, 0),      # W < 16 -> 16x8 patch
    (1, 24, 16, 256, 512, 1, 256), # three pixel tiles x two n-tiles
    (3, 8, 16, 512, 256, 2, 256),  # one n-tile, three pixel tiles
    (1, 24, 24, 256, 256, 3, 0),   # dil 3: not a halo-A pair (the halo stages would not fit), pairs with one A tile per tap
    (1, 20, 12, 128, 128, 3, 0),   # same, 128 wide (two k-chunks per stage)
    (3, 16, 8, 256, 256, 1, 0),    # three 8x16 pixel tiles: the pair's dead tile next to the staged TMA stores
    (1, 40, 24, 64, 256, 1, 0),    # one channel chunk per halo-A tile
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("nsplit", [1, 2])
def test_conv_store_matches_oracle(dev, case, nsplit):
    from can_b200 import ops
    N, H, W, Cin, Cout, dil, bn = case
    g = torch.Generator().manual_seed(3)
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * (2.0 / (9 * Cin)) ** 0.5
    b = torch.randn(Cout, generator=g) * 0.1
    xa = ops.act_from_nchw(x.to(dev), nsplit)
    xr = xa.to_nchw().cpu()           # what the kernel sees: bf16 (or hi+lo) rounding of x
    wp = ops.pack_conv_weight(w.to(dev), nsplit)
    wr = _bf16(w) if nsplit == 1 else _bf16(w) + _bf16(w - _bf16(w))
    ref = O.relu(O.conv3x3(xr.double().numpy(), wr.double().numpy(), b.double().numpy(), dil))
    out = ops.new_act(nsplit, N, H, W, Cout, dev, zero=True)
    ops.conv_tc(xa, wp, Cout, dil=dil, mode=ops.MODE_STORE, relu=True, bias=b.to(dev), out=out, bn=bn)
    got = out.to_nchw().cpu().double().numpy()
    rel = np.linalg.norm(got - ref) / np.linalg.norm(ref)
    # same operands, fp32 accumulation: the only error left is the output rounding (bf16: 2^-9 rms)
    assert rel < (3e-3 if nsplit == 1 else 3e-5), rel


def test_single_halo_tile_variant_forward(dev, monkeypatch):
    """CAN_B200_HALO1=2 (one halo tile for all nine taps): store and max-pool epilogues on 8-pixel-wide patches."""
    monkeypatch.setenv("CAN_B200_HALO1", "2")
    for case in [(1, 16, 16, 64, 64, 1, 0), (1, 16, 8, 64, 64, 1, 0), (2, 24, 40, 64, 128, 1, 0), (1, 13, 17, 128, 64, 2, 0)]:
        test_conv_store_matches_oracle(dev, case, 1)
    test_conv_pool_and_argmax(dev, 1)


@pytest.mark.parametrize("nsplit", [1, 2])
def test_conv_pool_and_argmax(dev, nsplit):
    from can_b200 import ops
    N, H, W, Cin, Cout = 2, 26, 34, 64, 128
    g = torch.Generator().manual_seed(5)
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * (2.0 / (9 * Cin)) ** 0.5
    b = torch.randn(Cout, generator=g) * 0.1
    xa = ops.act_from_nchw(x.to(dev), nsplit)
    wp = ops.pack_conv_weight(w.to(dev), nsplit)
    full = ops.new_act(nsplit, N, H, W, Cout, dev)
    ops.conv_tc(xa, wp, Cout, mode=ops.MODE_STORE, relu=True, bias=b.to(dev), out=full)
    pooled = ops.new_act(nsplit, N, H // 2, W // 2, Cout, dev, zero=True)
    arg = torch.full((N, H // 2, W // 2, Cout), 9, device=dev, dtype=torch.uint8)
    ops.conv_tc(xa, wp, Cout, mode=ops.MODE_POOL, bias=b.to(dev), out=pooled, arg=arg)
    f = full.to_nchw().cpu().numpy()
    ref = O.maxpool2x2(f)
    got = pooled.to_nchw().cpu().numpy()
    assert np.array_equal(got, ref)  # identical rounding: bit-exact against the un-fused path
    code = arg.permute(0, 3, 1, 2).cpu().numpy().astype(np.int64)
    assert code.max() <= 3
    n_i, c_i, y_i, x_i = np.indices(code.shape)
    picked = f[n_i, c_i, 2 * y_i + code // 2, 2 * x_i + code % 2]
    assert np.array_equal(picked, ref)
    # first maximum in scan order (PyTorch's MaxPool2d tie-break): no earlier slot holds the max
    win = f[:, :, :H // 2 * 2, :W // 2 * 2].reshape(N, Cout, H // 2, 2, W // 2, 2).transpose(0, 1, 2, 4, 3, 5)
    win = win.reshape(N, Cout, H // 2, W // 2, 4)
    first = (win == ref[..., None]).argmax(-1)
    assert np.array_equal(first, code)


def test_conv_head_matches_oracle(dev):
    from can_b200 import ops
    N, H, W, Cin, Cout = 1, 13, 17, 128, 64
    g = torch.Generator().manual_seed(9)
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * (2.0 / (9 * Cin)) ** 0.5
    b = torch.randn(Cout, generator=g) * 0.1
    hw = torch.randn(1, 64, 1, 1, generator=g) * 0.2
    hb = torch.tensor([0.3])
    for nsplit in (1, 2):
        xa = ops.act_from_nchw(x.to(dev), nsplit)
        xr = xa.to_nchw().cpu()
        wr = _bf16(w) if nsplit == 1 else _bf16(w) + _bf16(w - _bf16(w))
        y = O.relu(O.conv3x3(xr.double().numpy(), wr.double().numpy(), b.double().numpy(), 2))
        ref = O.conv1x1(y, hw.double().numpy(), hb.double().numpy())
        out = ops.new_act(nsplit, N, H, W, Cout, dev)
        dens = torch.zeros(N, 1, H, W, device=dev)
        ops.conv_tc(xa, ops.pack_conv_weight(w.to(dev), nsplit), Cout, dil=2, mode=ops.MODE_HEAD, bias=b.to(dev),
                    out=out, head_w=hw.reshape(64).to(dev), head_b=hb.to(dev), head_out=dens)
        rel = np.linalg.norm(dens.cpu().double().numpy() - ref) / np.linalg.norm(ref)
        assert rel < 2e-5, rel


# ---------------------------------------------------------------- K1: pooling + tables
@pytest.mark.parametrize("hw", [(8, 12), (13, 17), (48, 64), (6, 10)])
def test_ctx_pool_matches_oracle(dev, hw):
    from can_b200 import ops
    h, w = hw
    g = torch.Generator().manual_seed(11)
    fv = torch.rand(2, 512, h, w, generator=g)
    w1 = [torch.randn(512, 512, 1, 1, generator=g) * 0.05 for _ in range(4)]
    w2 = [torch.randn(512, 512, 1, 1, generator=g) * 0.05 for _ in range(4)]
    for nsplit in (1, 2):
        fa = ops.act_from_nchw(fv.to(dev), nsplit)
        fr = fa.to_nchw().cpu().double().numpy()
        pooled, tabA, tabB = ops.ctx_pool(fa, ops.stack_1x1([t.to(dev) for t in w1]),
                                          ops.stack_1x1([t.to(dev) for t in w2]))
        off = 0
        for j, s in enumerate(O.SCALES):
            ref = O.adaptive_avg_pool(fr, s)                        # [N,512,s,s]
            got = pooled[:, off:off + s * s].cpu().double().numpy()  # [N,s*s,512]
            ref_p = ref.reshape(2, 512, s * s).transpose(0, 2, 1)
            assert np.abs(got - ref_p).max() < 1e-5
            a_ref = O.conv1x1(ref, w1[j].double().numpy())
            b_ref = O.conv1x1(a_ref, w2[j].double().numpy())
            ga = tabA[:, off:off + s * s].cpu().double().numpy()
            gb = tabB[:, off:off + s * s].cpu().double().numpy()
            assert np.abs(ga - a_ref.reshape(2, 512, s * s).transpose(0, 2, 1)).max() < 1e-4
            assert np.abs(gb - b_ref.reshape(2, 512, s * s).transpose(0, 2, 1)).max() < 1e-4
            off += s * s


# ---------------------------------------------------------------- whole forward vs golden vectors
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(CASES))
def test_forward_matches_reference_golden(dev, golden, name, precision):
    regime, n, h, w = CASES[name]
    m = _model(regime, precision, dev)
    x = harness.make_input(n, h, w).to(dev)
    with torch.no_grad():
        out = m(x)
    assert out.shape == (n, 1, h // 8, w // 8) and out.dtype == torch.float32
    ref = torch.from_numpy(golden[f"{name}/f64/out"])
    tol = TOL[precision]
    if regime == "ref" and precision == "bf16":
        pytest.skip("reference init gives ~1e-10 outputs: bf16 relative error is meaningless (SURVEY 8c)")
    assert harness.rel_l2(out.cpu(), ref) < tol["l2"]
    assert harness.rel_count(out.cpu(), ref) < tol["count"]


def test_ctx2_direct_store_variant_matches_reference(dev, golden, monkeypatch):
    """CAN_B200_CTX2STAGE=0: the contextual-module GEMM writes its fused features with per-lane stores instead of the
    staged TMA store; same golden outputs."""
    monkeypatch.setenv("CAN_B200_CTX2STAGE", "0")
    for name in [n for n in CASES if CASES[n][0] != "ref"]:   # ("ref" is skipped for bf16, see above)
        test_forward_matches_reference_golden(dev, golden, name, "bf16")


def test_fv_and_fi_match_reference(dev, golden):
    from can_b200 import engine
    name = "he_1x64x96"
    regime, n, h, w = CASES[name]
    m = _model(regime, "fp32", dev)
    with torch.no_grad():
        _, saved = engine.forward(m, harness.make_input(n, h, w).to(dev), m.nsplit, save=True)
    fv = saved["cat"].view(512, 0).to_nchw().cpu()
    fi = saved["cat"].view(512, 512).to_nchw().cpu()
    assert harness.rel_l2(fv, torch.from_numpy(golden[f"{name}/fv"])) < 5e-5
    assert harness.rel_l2(fi, torch.from_numpy(golden[f"{name}/fi"])) < 5e-5


def test_forward_is_batch_consistent_at_full_size(dev):
    """Size-independent property at BASELINE's 768x1024: images are independent (no BN), so a
    batch of two copies gives two bit-identical maps equal to the single-image result."""
    m = _model("he", "bf16", dev)
    x = harness.make_input(1, 768, 1024).to(dev)
    with torch.no_grad():
        a = m(x)
        b = m(torch.cat([x, x], 0))
    assert a.shape == (1, 1, 96, 128)
    assert torch.equal(b[0], b[1]) and torch.equal(a[0], b[0])
    assert torch.isfinite(a).all()


def test_no_grad_and_eval_modes_agree(dev):
    m = _model("he", "bf16", dev)
    x = harness.make_input(1, 64, 96).to(dev)
    with torch.no_grad():
        a = m(x)
        m.eval()
        b = m(x)
    assert torch.equal(a, b)


# ---------------------------------------------------------------- loss / count / SGD
def test_mse_count_matches_oracle(dev):
    from can_b200 import ops
    g = torch.Generator().manual_seed(2)
    for n, h, w in [(1, 8, 12), (3, 96, 128), (16, 96, 128)]:
        et = torch.randn(n, 1, h, w, generator=g)
        gt = torch.rand(n, 1, h, w, generator=g) * 0.05
        for _ in range(2):  # the workspace counter must rearm itself
            loss, ce, cg, grad = ops.mse_count(et.to(dev), gt.to(dev), want_grad=True)
            assert float(loss) == pytest.approx(O.mse_sum(et.numpy(), gt.numpy()), rel=2e-6)
            assert np.allclose(ce.cpu().numpy(), O.counts(et.numpy()), rtol=1e-5, atol=1e-3)
            assert np.allclose(cg.cpu().numpy(), O.counts(gt.numpy()), rtol=1e-5, atol=1e-3)
            assert np.allclose(grad.cpu().numpy(), O.mse_sum_grad(et.numpy(), gt.numpy()), rtol=1e-6, atol=1e-7)


def test_sgd_matches_oracle(dev):
    from can_b200 import ops
    g = torch.Generator().manual_seed(4)
    n = 1000003
    p = torch.randn(n, generator=g)
    pn, buf_n = p.double().numpy().copy(), None
    pd, bd = p.to(dev), torch.zeros(n, device=dev)
    for step in range(3):
        gr = torch.randn(n, generator=g)
        ops.sgd_momentum_step(pd, gr.to(dev), bd, 1e-3, 0.95, 1.0, first=(step == 0))
        pn, buf_n = O.sgd_momentum_step(pn, gr.double().numpy(), buf_n, 1e-3, 0.95)
    assert np.allclose(pd.cpu().numpy(), pn, rtol=1e-6, atol=1e-7)
    assert np.allclose(bd.cpu().numpy(), buf_n, rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("cout", [64, 128])
def test_conv_store_into_a_channel_window(dev, cout):
    """The staged (TMA-store) epilogue writing a channel window [coff, coff + Cout) of a wider NHWC buffer, ragged tiles:
    same values as a store into a dense tensor, and the neighbouring channels untouched."""
    from can_b200 import ops
    N, H, W, Cin = 2, 19, 37, 64
    g = torch.Generator().manual_seed(8)
    x = ops.act_from_nchw(torch.randn(N, Cin, H, W, generator=g).to(dev), 1)
    w = (torch.randn(cout, Cin, 3, 3, generator=g) * 0.06).to(dev)
    b = (torch.randn(cout, generator=g) * 0.1).to(dev)
    wp = ops.pack_conv_weight(w, 1)
    dense = ops.new_act(1, N, H, W, cout, dev, zero=True)
    ops.conv_tc(x, wp, cout, mode=ops.MODE_STORE, relu=True, bias=b, out=dense)
    wide = ops.new_act(1, N, H, W, cout + 128, dev, zero=True)
    wide.buf.fill_(7.0)
    ops.conv_tc(x, wp, cout, mode=ops.MODE_STORE, relu=True, bias=b, out=wide.view(cout, 64))
    got = wide.buf.view(N, H, W, cout + 128)
    assert torch.equal(got[..., 64:64 + cout], dense.buf.view(N, H, W, cout))
    assert bool((got[..., :64] == 7.0).all()) and bool((got[..., 64 + cout:] == 7.0).all())


@pytest.mark.gpu
def test_inference_graph_replay_equals_eager_and_follows_weight_changes(dev):
    """no_grad calls of one shape replay a CUDA graph from the third call on: same output as the eager pass, results do
    not alias each other, and a parameter change (load_state_dict / in-place update + cache invalidation) is picked up."""
    m = _model("he", "bf16", dev)
    x = harness.make_input(2, 64, 96).to(dev)
    x2 = harness.make_input(2, 64, 96, seed=77).to(dev)
    with torch.no_grad():
        outs = [m(x) for _ in range(4)]           # eager, eager, capture + replay, replay
        assert all(torch.equal(outs[0], o) for o in outs[1:])
        o2 = m(x2)                                # replay with other data
        assert not torch.equal(o2, outs[0]) and torch.equal(outs[3], outs[0])   # earlier results were not overwritten
        m.load_state_dict(harness.make_state_dict("default"))
        ref = _model("default", "bf16", dev)(x)
        assert torch.equal(m(x), ref)             # new weights -> graph dropped, eager again
        with torch.no_grad():
            m.backend[10].bias.data.add_(0.5)     # in-place change without a version bump on .data ...
        m._wcache.invalidate()                    # ... announced the way FusedTrainer does
        assert not torch.equal(m(x), ref)
