"""Execution plan of the CANNet step on the sm_100a kernels.

``forward`` runs model/CANNet.py:39-91 of the reference as 17 tcgen05 implicit-GEMM
launches plus the small HBM-bound kernels; ``backward`` (autograd of the same graph,
utils/train_eval_utils.py:38) reuses the saved NHWC activations: per conv one dgrad launch
(the same implicit-GEMM kernel on transposed/flipped weights with the ReLU mask or max-unpool
fused) and one wgrad launch (MN-major tcgen05 GEMM over pixels).  Packed weights are cached
per parameter version, so a training loop re-packs after every optimizer step and an
inference loop packs once.

Parameter order everywhere = ``model.parameters()`` = the reference's state_dict order:
frontend (w,b) x10, backend (w,b) x6, output_layer (w,b), conv{1,2,3,6}_{1,2}.weight.
"""
from __future__ import annotations

import os

import torch

from . import ops
from .utils import nvtx
from .ops import (MODE_CTX, MODE_CTX2, MODE_CTX2_BWD, MODE_CTX_BWD, MODE_DFV, MODE_HEAD, MODE_POOL, MODE_STORE,
                  MODE_UNPOOL, Act)

FRONTEND_CONVS = (0, 2, 5, 7, 10, 12, 14, 17, 19, 21)  # Sequential indices, CANNet.py:104-121
POOLED_LAYERS = (1, 3, 6)                               # frontend convs followed by MaxPool2d(2,2)
BACKEND_CONVS = (0, 2, 4, 6, 8, 10)
SCALES = (1, 2, 3, 6)
IDX_BACKEND, IDX_HEAD, IDX_CTX = 20, 32, 34             # offsets into the parameter list


class WeightCache:
    """Packed (kernel-layout) copies of the module's parameters, rebuilt when a parameter's
    storage or version counter changes (load_state_dict, optimizer.step, .to(device))."""

    def __init__(self):
        self._store = {}
        self.epoch = 0      # bumped by invalidate(): lets graph replays notice in-place parameter updates

    def get(self, key, params, build):
        tag = tuple((p.data_ptr(), p._version, p.device) for p in params)
        hit = self._store.get(key)
        if hit is None or hit[0] != tag:
            with torch.no_grad():
                hit = (tag, build())
            self._store[key] = hit
        return hit[1]

    def put(self, key, params, value):
        """Store a value built elsewhere (a multi-tensor pack) under the tag `get` would compute."""
        self._store[key] = (tuple((p.data_ptr(), p._version, p.device) for p in params), value)

    def fresh(self, key, params):
        hit = self._store.get(key)
        return hit is not None and hit[0] == tuple((p.data_ptr(), p._version, p.device) for p in params)

    def invalidate(self):
        """Drop every entry derived from parameters (after an in-place arena update, which does
        not bump version counters); shape-only entries (empty tag) are kept."""
        self._store = {k: v for k, v in self._store.items() if v[0] == ()}
        self.epoch += 1


def _packed(cache, tag, li, wt, nsplit, both):
    """(fwd pack[, dgrad pack]) of one conv weight; a training step builds both in one pass."""
    if both:
        return cache.get((tag, li, nsplit, "both"), (wt,), lambda: ops.pack_conv_weight_both(wt, nsplit))
    return (cache.get((tag, li, nsplit), (wt,), lambda: ops.pack_conv_weight(wt, nsplit)),)


def prepack(model, nsplit):
    """Training step: the forward and data-gradient packings of all fifteen 3x3 conv weights in ONE launch (they are
    all stale after an optimizer update; fifteen small launches otherwise)."""
    cache: WeightCache = model._wcache
    fr, bk = _conv_params(model)
    todo = [(("f", li, nsplit, "both"), fr[li][0]) for li in range(1, 10)] + \
           [(("b", li, nsplit, "both"), bk[li][0]) for li in range(6)]
    todo = [(k, w) for k, w in todo if not cache.fresh(k, (w,)) and w.is_contiguous() and w.dtype == torch.float32]
    if len(todo) < 2:
        return
    with torch.no_grad():
        packs = ops.pack_conv_weights_multi([w for _, w in todo], nsplit)
    for (k, w), pk in zip(todo, packs):
        cache.put(k, (w,), pk)


def _conv_params(model):
    fr = [(model.frontend[i].weight, model.frontend[i].bias) for i in FRONTEND_CONVS]
    bk = [(model.backend[i].weight, model.backend[i].bias) for i in BACKEND_CONVS]
    return fr, bk


def _ctx_weights(model):
    w1s = [getattr(model, f"conv{s}_1").weight for s in SCALES]
    w2s = [getattr(model, f"conv{s}_2").weight for s in SCALES]
    return w1s, w2s


def forward(model, x: torch.Tensor, nsplit: int, save: bool):
    """Returns (density [N,1,h,w] fp32, saved dict or None)."""
    cache: WeightCache = model._wcache
    dev = x.device
    N, _, H, W = x.shape
    x = x.contiguous().float()
    fr, bk = _conv_params(model)
    saved = {"x": x} if save else None
    if save:
        prepack(model, nsplit)

    # ---- frontend (CANNet.py:40) -------------------------------------------------
    nvtx.push("can.fwd.frontend")
    w0, b0 = fr[0]
    col = None
    first_mode = os.environ.get("CAN_B200_FIRST_TC", "2") if nsplit == 1 else "0"
    if first_mode == "2":
        # bf16 path: conv1_1 as a K = 32 GEMM whose 27-wide patch rows are rebuilt in shared memory from the
        # fp32 image (no im2col buffer in HBM); backward does the same for the weight gradient
        w0p = cache.get(("f0", 1), (w0,), lambda: ops.pack_first_weight(w0))
        a = ops.new_act(1, N, H, W, 64, dev)
        ops.conv_first_tc(x, w0p, b0.detach().float().contiguous(), a)
    elif first_mode == "1":
        # K = 27 padded to 64 through an im2col of the image (1x1 GEMM); the buffer is reused by backward
        col = ops.im2col_first(x, 1)
        w0p = cache.get(("f0", 1), (w0,), lambda: ops.pack_first_weight(w0))
        a = ops.new_act(1, N, H, W, 64, dev)
        ops.conv_tc(col, w0p, 64, taps=1, mode=MODE_STORE, relu=True, bias=b0.detach().float(), out=a)
    else:
        a = ops.conv_first(x, w0.detach().float().contiguous(), b0.detach().float().contiguous(), nsplit)
    acts = [a]
    args = {}
    h, w = H, W
    cat = None
    for li in range(1, 10):
        wt, bs = fr[li]
        cout = wt.shape[0]
        wp = _packed(cache, "f", li, wt, nsplit, save)[0]
        bias = bs.detach().float()
        if li in POOLED_LAYERS:
            h, w = h // 2, w // 2
            out = ops.new_act(nsplit, N, h, w, cout, dev)
            arg = torch.empty(N, h, w, cout, device=dev, dtype=torch.uint8) if save else None
            ops.conv_tc(a, wp, cout, mode=MODE_POOL, bias=bias, out=out, arg=arg)
            args[li] = arg
        elif li == 9:
            # conv4_3 writes fv straight into channels [0,512) of the backend input (no torch.cat)
            cat = ops.new_act(nsplit, N, h, w, 1024, dev)
            out = cat.view(512, 0)
            ops.conv_tc(a, wp, cout, mode=MODE_STORE, relu=True, bias=bias, out=out)
        else:
            out = ops.new_act(nsplit, N, h, w, cout, dev)
            ops.conv_tc(a, wp, cout, mode=MODE_STORE, relu=True, bias=bias, out=out)
        a = out
        acts.append(a)
    fv = a
    nvtx.pop()

    # ---- contextual module (CANNet.py:42-87) --------------------------------------
    nvtx.push("can.fwd.context")
    w1s, w2s = _ctx_weights(model)
    # (inside a FusedTrainer arena the four tables of a kind lie back to back: a view, no copy)
    def stacked(ws):
        v = _stacked_view([w_.detach() for w_ in ws])
        return v if v is not None else ops.stack_1x1(ws)
    w1 = cache.get(("w1",), w1s, lambda: stacked(w1s))
    w2 = cache.get(("w2",), w2s, lambda: stacked(w2s))
    w2stack = cache.get(("w2stack", nsplit), w2s, lambda: ops.pack_ctx_w2(w2s, nsplit))
    pooled, tabA, tabB = ops.ctx_pool(fv, w1, w2)
    fi = cat.view(512, 512)
    if nsplit == 1 and os.environ.get("CAN_B200_CTX2", "1") != "0":
        # bf16 path: the bilinear interpolation of both tables is two extra k-blocks of the same GEMM
        wint1 = cache.get(("wint", 1, h, w), (), lambda: ops.ctx_wint(1, 1, h, w, dev))
        ttab = ops.ctx_ttab(tabA, tabB)
        ops.conv_tc(fv, w2stack, 2048, taps=1, mode=MODE_CTX2, out=fi, ctx_wint=wint1.buf, ctx_ttab=ttab)
    else:
        ttab = None
        ops.conv_tc(fv, w2stack, 2048, taps=1, mode=MODE_CTX, out=fi, tabA=tabA, tabB=tabB)

    nvtx.pop()

    # ---- backend + head (CANNet.py:89-90) -----------------------------------------
    nvtx.push("can.fwd.backend")
    a = cat.view(1024, 0)
    bacts = [a]
    density = torch.empty(N, 1, h, w, device=dev, dtype=torch.float32)
    for li in range(6):
        wt, bs = bk[li]
        cout = wt.shape[0]
        wp = _packed(cache, "b", li, wt, nsplit, save)[0]
        bias = bs.detach().float()
        out = ops.new_act(nsplit, N, h, w, cout, dev)
        if li == 5:
            ow = model.output_layer.weight.detach().float().reshape(64).contiguous()
            ob = model.output_layer.bias.detach().float().contiguous()
            ops.conv_tc(a, wp, cout, dil=2, mode=MODE_HEAD, bias=bias, out=out, head_w=ow, head_b=ob,
                        head_out=density)
        else:
            ops.conv_tc(a, wp, cout, dil=2, mode=MODE_STORE, relu=True, bias=bias, out=out)
        a = out
        bacts.append(a)
    nvtx.pop()
    if save:
        saved.update(acts=acts, args=args, cat=cat, bacts=bacts, pooled=pooled, tabA=tabA, tabB=tabB, ttab=ttab, col=col,
                     nsplit=nsplit)
    return density, saved


def new_grads(model):
    """Gradient tensors in parameter order (biases and the head start at zero: they are
    accumulated with atomics; weights are fully overwritten by the wgrad reduction)."""
    out = []
    for p in model.parameters():
        out.append(torch.zeros_like(p) if p.dim() == 1 or p.shape[0] == 1 else torch.empty_like(p))
    return out


def backward(model, saved, grad_out: torch.Tensor, grads=None, on_ready=None):
    """Gradients of every parameter given dL/d(density).  ``grads`` (optional) are the fp32
    tensors to fill, in parameter order, with bias / head entries already zeroed.
    ``on_ready(tag)`` is called after the gradients of 'backend', 'ctx' and 'frontend' have been
    enqueued (used to start the bucketed all-reduce while the rest of backward still runs)."""
    gen = backward_steps(model, saved, grad_out, grads)
    while True:
        try:
            tag = next(gen)
        except StopIteration as done:
            return done.value
        if on_ready:
            on_ready(tag)


def _stacked_view(gs):
    """Four contiguous fp32 [512,512,1,1] tensors lying back to back in one storage (16-byte aligned) as one
    [4,512,512] view of that storage, else None."""
    g0 = gs[0]
    n = g0.numel()
    ok = g0.data_ptr() % 16 == 0 and all(
        g.is_contiguous() and g.dtype == torch.float32 and g.untyped_storage().data_ptr() == g0.untyped_storage().data_ptr()
        and g.storage_offset() == g0.storage_offset() + j * n for j, g in enumerate(gs))
    return g0.as_strided((len(gs), 512, 512), (n, 512, 1)) if ok else None


class _SideStream:
    """Weight gradients of a layer and the data gradient that feeds the next layer are independent: the wgrad
    kernels go to a second stream so that the tail of one persistent kernel (e.g. 768 tiles on 148 SMs = a sixth,
    19 %-full wave) overlaps with the other.  Works eagerly and under CUDA-graph capture (fork/join by events)."""
    _streams = {}

    def __init__(self, dev, enabled):
        self.enabled = enabled
        self.keep = []
        if enabled:
            self.main = torch.cuda.current_stream(dev)
            self.side = _SideStream._streams.get(dev)
            if self.side is None:
                self.side = _SideStream._streams[dev] = torch.cuda.Stream(dev)

    def run(self, fn, *tensors):
        """fn() on the side stream, after everything enqueued so far on the main stream."""
        if not self.enabled:
            fn()
            return
        self.side.wait_stream(self.main)
        with torch.cuda.stream(self.side):
            fn()
        self.keep.extend(tensors)   # operands stay alive (and their memory unreused) until the join

    def join(self):
        if self.enabled:
            self.main.wait_stream(self.side)
        self.keep.clear()


def backward_steps(model, saved, grad_out: torch.Tensor, grads=None):
    """``backward`` as a generator: yields 'backend', 'ctx', 'frontend_hi' (conv4_x: 5.9 of the frontend's 7.6 M
    gradients, ready ~2 ms before the rest) and 'frontend' right after the gradients of that bucket have been enqueued,
    returns the gradient list.  (The data-parallel trainer captures each section as its own CUDA graph and issues the
    NCCL all-reduces between the replays.)"""
    cache: WeightCache = model._wcache
    nsplit = saved["nsplit"]
    fr, bk = _conv_params(model)
    acts, args, cat, bacts = saved["acts"], saved["args"], saved["cat"], saved["bacts"]
    dev = cat.buf.device
    N, h, w = cat.N, cat.H, cat.W
    if grads is None:
        grads = new_grads(model)
    g = grad_out.contiguous().float()
    # (measured on the power-capped B200s of this pool, three interleaved 40-step runs each: 8.40 -> 8.33 ms per step,
    # profiles/r02_side_ab.txt; CAN_B200_SIDE=0 keeps everything on one stream)
    side = _SideStream(dev, os.environ.get("CAN_B200_SIDE", "1") == "1")

    # ---- head + backend -----------------------------------------------------------
    nvtx.push("can.bwd.backend")
    hw = model.output_layer.weight.detach().float().reshape(64).contiguous()
    dz = ops.head_bwd(g, bacts[6], hw, grads[IDX_HEAD], grads[IDX_HEAD + 1])
    dcat = None
    for li in range(5, -1, -1):
        wt = bk[li][0]
        cin = wt.shape[1]
        X = bacts[li]
        side.run(lambda dz=dz, X=X, li=li: ops.conv_wgrad(dz, X, grads[IDX_BACKEND + 2 * li], dil=2), dz, X)
        if li == 5:
            ops.bias_grad(dz, grads[IDX_BACKEND + 2 * li + 1])   # dz of the last conv comes from head_bwd
        wd = _packed(cache, "b", li, wt, nsplit, True)[1]
        ndz = ops.new_act(nsplit, N, h, w, cin, dev)
        if li > 0:
            # the dgrad epilogue also accumulates the bias gradient of the layer whose dZ it writes
            ops.conv_tc(dz, wd, cin, dil=2, mode=MODE_STORE, relu=False, out=ndz, mask=X,
                        db=grads[IDX_BACKEND + 2 * (li - 1) + 1])
            dz = ndz
        else:
            ops.conv_tc(dz, wd, cin, dil=2, mode=MODE_STORE, relu=False, out=ndz)
            dcat = ndz
    side.join()
    nvtx.pop()
    yield "backend"
    nvtx.push("can.bwd.context")

    # ---- contextual module ----------------------------------------------------------
    w1s, w2s = _ctx_weights(model)
    fv = cat.view(512, 0)
    w2stack = cache.get(("w2stack", nsplit), w2s, lambda: ops.pack_ctx_w2(w2s, nsplit))
    wt = cache.get(("w12t",), w1s + w2s, lambda: ops.transpose_stack(w1s + w2s))   # eight W^T tables, one launch
    w1t, w2t = wt[:4], wt[4:]
    D = ops.new_act(nsplit, N, h, w, 4096, dev)   # [dz_1..dz_6 | direct ds_1..ds_6], 512 channels each
    if saved.get("ttab") is not None:
        wint1 = cache.get(("wint", 1, h, w), (), lambda: ops.ctx_wint(1, 1, h, w, dev))
        ops.conv_tc(fv, w2stack, 2048, taps=1, mode=MODE_CTX2_BWD, out=D, mask=dcat.view(512, 512), ctx_wint=wint1.buf,
                    ctx_ttab=saved["ttab"])
    else:
        ops.conv_tc(fv, w2stack, 2048, taps=1, mode=MODE_CTX_BWD, out=D, mask=dcat.view(512, 512), tabA=saved["tabA"],
                    tabB=saved["tabB"])
    wint = cache.get(("wint", nsplit, h, w), (), lambda: ops.ctx_wint(nsplit, 1, h, w, dev))
    if nsplit == 1:
        T = ops.ctx_ut(wint, D)                      # U^T [dz | ds], [64 pos][N][4096]: one launch for the batch
        t_ns, t_ps = 4096, N * 4096
    else:
        T = torch.empty(N, 64, 4096, device=dev, dtype=torch.float32)   # per image
        for n in range(N):
            ops.conv_wgrad(wint, D.image(n), T[n], taps=1)
        t_ns, t_ps = 64 * 4096, 4096
    # da_j = U^T ds_j + W_j2^T (U^T dz_j);  dpool_j = W_j1^T da_j
    da = ops.ctx_pos_gemv(w2t, T, t_ns, t_ps, 512, T.view(-1)[2048:], N)
    dpool = ops.ctx_pos_gemv(w1t, da, 50 * 512, 512, 0, None, N)
    # the four conv{S}_1 (conv{S}_2) gradients as one [4,512,512] tensor: written in place when the caller's gradient
    # tensors lie back to back (the trainer's arena does), through a temporary + four copies otherwise
    dW1 = _stacked_view([grads[IDX_CTX + 2 * j] for j in range(4)])
    dW2 = _stacked_view([grads[IDX_CTX + 2 * j + 1] for j in range(4)])
    direct = dW1 is not None and dW2 is not None
    if not direct:
        dW1 = torch.empty(4, 512, 512, device=dev, dtype=torch.float32)
        dW2 = torch.empty(4, 512, 512, device=dev, dtype=torch.float32)
    ops.ctx_outer(da, 50 * 512, 512, 0, saved["pooled"], N, dW1)          # dW_j1 = sum da_j pool_j^T
    ops.ctx_outer(T, t_ns, t_ps, 512, saved["tabA"], N, dW2)              # dW_j2 = sum (U^T dz_j) a_j^T ...
    # ... - dz_j fv^T: the four scales share fv, so one [2048 x 512] weight-gradient GEMM over all of them
    ops.conv_wgrad(D.view(2048, 0), fv, dW2.view(2048, 512), taps=1, accumulate=True, alpha=-1.0)
    if not direct:
        for j in range(4):
            grads[IDX_CTX + 2 * j].copy_(dW1[j].view(512, 512, 1, 1))
            grads[IDX_CTX + 2 * j + 1].copy_(dW2[j].view(512, 512, 1, 1))
    # dfv = dcat[:, :512] - sum_j W_j2^T dz_j + pool^T(dpool), gated by fv > 0  -> dZ of conv4_3
    w2d = cache.get(("w2d", nsplit), w2s, lambda: ops.pack_conv_weight(
        torch.cat([wj.detach().float().reshape(512, 512).t() for wj in w2s], dim=1).reshape(512, 2048, 1, 1), nsplit))
    dz = ops.new_act(nsplit, N, h, w, 512, dev)
    pint = dtab = None
    if nsplit == 1:
        # bf16 path: the pooling-transpose term as one more k-block of the same GEMM (the kernel takes it when the launch
        # runs as 256-wide CTA pairs, else it adds the table rows in its epilogue as the fp32 path does)
        pint = cache.get(("pint", h, w), (), lambda: ops.ctx_pint(h, w, dev))
        dtab = ops.ctx_dtab(dpool)
    ops.conv_tc(D.view(2048, 0), w2d, 512, taps=1, mode=MODE_DFV, relu=False, out=dz, mask=fv, add=dcat.view(512, 0),
                tabA=dpool, db=grads[2 * 9 + 1], ctx_wint=pint, ctx_ttab=dtab)
    nvtx.pop()
    yield "ctx"
    nvtx.push("can.bwd.frontend")

    # ---- frontend -------------------------------------------------------------------
    for li in range(9, 0, -1):
        wt = fr[li][0]
        cin = wt.shape[1]
        X = acts[li - 1]
        side.run(lambda dz=dz, X=X, li=li: ops.conv_wgrad(dz, X, grads[2 * li], dil=1), dz, X)
        wd = _packed(cache, "f", li, wt, nsplit, True)[1]
        dbp = grads[2 * (li - 1) + 1]   # bias gradient of layer li-1, fused into the dgrad epilogue below
        if (li - 1) in POOLED_LAYERS:
            pre = acts[li - 2]  # the pooled layer's input has its pre-pool resolution
            OH, OW = pre.H, pre.W
            ndz = ops.new_act(nsplit, N, OH, OW, cin, dev, zero=bool((OH | OW) & 1))
            ops.conv_tc(dz, wd, cin, mode=MODE_UNPOOL, relu=False, out=ndz, mask=X, arg=args[li - 1], OH=OH, OW=OW,
                        db=dbp)
        else:
            ndz = ops.new_act(nsplit, N, X.H, X.W, cin, dev)
            ops.conv_tc(dz, wd, cin, mode=MODE_STORE, relu=False, out=ndz, mask=X, db=dbp)
        dz = ndz
        if li == 7:
            side.join()
            nvtx.pop()
            yield "frontend_hi"   # weights of conv4_1..conv4_3 and the biases next to them are final
            nvtx.push("can.bwd.frontend")
    # first conv (3 -> 64): K = 27 padded to 64 through an im2col of the image
    col = saved.get("col")
    if nsplit == 1 and col is None and grads[0].is_contiguous() and os.environ.get("CAN_B200_FIRST_TC", "2") == "2":
        ops.conv_first_wgrad(dz, saved["x"], grads[0])   # straight from the image, into the parameter's layout
    else:
        side.join()   # this weight gradient runs on the main stream and shares the split-K workspace with the side stream's
        if col is None:
            col = ops.im2col_first(saved["x"], nsplit)
        g0 = torch.empty(64, 64, device=dev, dtype=torch.float32)
        ops.conv_wgrad(dz, col, g0, taps=1)
        grads[0].copy_(g0[:, :27].reshape(64, 3, 3, 3))
    # (the bias gradient of conv1_1 was fused into conv1_2's dgrad)
    side.join()
    nvtx.pop()
    yield "frontend"
    return grads
