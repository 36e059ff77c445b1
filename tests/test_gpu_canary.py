"""Out-of-bounds writes, checked by hand (compute-sanitizer is closed on this GPU pool): every output of the main kernels
is a window inside a larger allocation pre-filled with a sentinel; after the launch the guard bands on both sides, and the
channels next to a channel window, must still hold the sentinel.  Ragged shapes (partial tiles, odd sizes) on purpose:
the TMA stores clip at the tensor-map bounds, the per-lane stores at their `valid` predicate."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SENT = 12345.0
GUARD = 4096  # elements on each side


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _guarded(shape, dtype, dev, fill=SENT):
    n = 1
    for s in shape:
        n *= s
    flat = torch.full((n + 2 * GUARD,), fill, dtype=dtype, device=dev)
    return flat, flat[GUARD:GUARD + n].view(*shape)


def _bands_intact(flat, fill=SENT):
    return bool((flat[:GUARD] == fill).all()) and bool((flat[-GUARD:] == fill).all())


@pytest.mark.parametrize("case", [(2, 19, 37, 64, 64, 1), (1, 13, 17, 128, 128, 1), (1, 11, 23, 256, 256, 2), (3, 9, 21, 512, 512, 1),
                                  (1, 17, 9, 64, 128, 2)])
def test_conv_store_pool_unpool_stay_inside_their_windows(dev, case):
    from can_b200 import ops
    from can_b200.ops import Act
    N, H, W, Cin, Cout, dil = case
    g = torch.Generator().manual_seed(1)
    x = ops.act_from_nchw(torch.randn(N, Cin, H, W, generator=g).to(dev), 1)
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) * 0.05).to(dev)
    b = torch.zeros(Cout, device=dev)
    wp = ops.pack_conv_weight(w, 1)
    # store into a channel window [64, 64 + Cout) of a wider pixel
    cs = Cout + 128
    flat, buf = _guarded((1, N, H, W, cs), torch.bfloat16, dev)
    out = Act(buf, Cout, 64)
    ops.conv_tc(x, wp, Cout, dil=dil, mode=ops.MODE_STORE, relu=True, bias=b, out=out)
    torch.cuda.synchronize()
    assert _bands_intact(flat)
    assert bool((buf[..., :64] == SENT).all()) and bool((buf[..., 64 + Cout:] == SENT).all())
    assert not bool((buf[..., 64:64 + Cout] == SENT).any())     # every element of the window was written
    if dil == 1:
        # fused 2x2 max-pool + arg-max codes (floor mode: the odd last row / column has no window)
        pf, pbuf = _guarded((1, N, H // 2, W // 2, Cout), torch.bfloat16, dev)
        af, abuf = _guarded((N, H // 2, W // 2, Cout), torch.uint8, dev, fill=77)
        ops.conv_tc(x, wp, Cout, mode=ops.MODE_POOL, bias=b, out=Act(pbuf, Cout, 0), arg=abuf)
        torch.cuda.synchronize()
        assert _bands_intact(pf) and _bands_intact(af, 77)
        assert not bool((pbuf == SENT).any()) and int(abuf.max()) <= 3
        # data gradient + max-unpool back onto the H x W grid from the pooled grid
        dz = ops.act_from_nchw(torch.randn(N, Cout, H // 2, W // 2, generator=g).to(dev), 1)
        wd = ops.pack_conv_weight((torch.randn(Cout, Cout, 3, 3, generator=g) * 0.05).to(dev), 1, dgrad=True)
        uf, ubuf = _guarded((1, N, H, W, Cout), torch.bfloat16, dev)
        ubuf.zero_()
        ops.conv_tc(dz, wd, Cout, mode=ops.MODE_UNPOOL, relu=False, out=Act(ubuf, Cout, 0), mask=Act(pbuf, Cout, 0), arg=abuf,
                    OH=H, OW=W)
        torch.cuda.synchronize()
        assert _bands_intact(uf)
        if H % 2:
            assert bool((ubuf[:, :, H - 1] == 0).all())     # the uncovered last row stays zero
        if W % 2:
            assert bool((ubuf[:, :, :, W - 1] == 0).all())


@pytest.mark.parametrize("case", [(2, 13, 21, 64, 64, 1), (1, 11, 19, 128, 256, 2), (1, 9, 17, 512, 512, 1)])
def test_wgrad_and_workspace_stay_inside(dev, case):
    from can_b200 import ops
    N, H, W, Cin, Cout, dil = case
    g = torch.Generator().manual_seed(2)
    x = ops.act_from_nchw(torch.relu(torch.randn(N, Cin, H, W, generator=g)).to(dev), 1)
    dz = ops.act_from_nchw(torch.randn(N, Cout, H, W, generator=g).to(dev), 1)
    gf, grad = _guarded((Cout, Cin, 3, 3), torch.float32, dev)
    ops.conv_wgrad(dz, x, grad, dil=dil)
    torch.cuda.synchronize()
    assert _bands_intact(gf) and not bool((grad == SENT).any())


@pytest.mark.parametrize("shape", [(2, 13, 17), (3, 48, 64), (1, 96, 128)])
def test_contextual_kernels_stay_inside(dev, shape):
    from can_b200 import ops
    from can_b200.ops import Act
    n, h, w = shape
    g = torch.Generator().manual_seed(3)
    flat, buf = _guarded((1, n, h, w, 1024), torch.bfloat16, dev)
    buf.copy_(torch.rand(1, n, h, w, 1024, generator=g).to(dev))
    buf[..., 512:] = SENT
    cat = Act(buf, 1024, 0)
    fv, fi = cat.view(512, 0), cat.view(512, 512)
    w1 = torch.randn(4, 512, 512, generator=g).to(dev) * 0.05
    w2l = [(torch.randn(512, 512, 1, 1, generator=g) * 0.05).to(dev) for _ in range(4)]
    w2 = torch.stack([t.reshape(512, 512) for t in w2l])
    before = buf[..., :512].clone()
    pooled, tabA, tabB = ops.ctx_pool(fv, w1, w2)
    ttab = ops.ctx_ttab(tabA, tabB)
    wint = ops.ctx_wint(1, 1, h, w, dev)
    ops.conv_tc(fv, ops.pack_ctx_w2(w2l, 1), 2048, taps=1, mode=ops.MODE_CTX2, out=fi, ctx_wint=wint.buf, ctx_ttab=ttab)
    torch.cuda.synchronize()
    assert _bands_intact(flat)
    assert torch.equal(buf[..., :512], before)                  # fv untouched
    assert not bool((buf[..., 512:] == SENT).any())             # fi fully written
    assert torch.isfinite(pooled).all() and torch.isfinite(tabA).all() and torch.isfinite(tabB).all()


def test_loss_and_sgd_stay_inside(dev):
    from can_b200 import ops
    g = torch.Generator().manual_seed(4)
    n, hw = 5, 37 * 41
    et = torch.randn(n, 1, 37, 41, generator=g).to(dev)
    gt = torch.rand(n, 1, 37, 41, generator=g).to(dev)
    loss, ce, cg, grad = ops.mse_count(et, gt, want_grad=True)
    assert grad.shape == et.shape and torch.isfinite(grad).all() and ce.numel() == n
    m = 100003
    pf, p = _guarded((m,), torch.float32, dev)
    bf, b = _guarded((m,), torch.float32, dev)
    p.normal_()
    b.zero_()
    ops.sgd_momentum_step(p, torch.randn(m, device=dev), b, 1e-3, 0.95, 1.0, first=True)
    torch.cuda.synchronize()
    assert _bands_intact(pf) and _bands_intact(bf)
