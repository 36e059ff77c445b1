"""Torch-tensor front ends of the C-ABI entry points (device memory + streams only).

``Act`` describes an activation in the kernels' layout: bf16 ``[plane][N][H][W][cs]``
where the tensor occupies channels ``[coff, coff + C)`` of the pixel stride ``cs``.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import torch

from ._lib import ConvArgs, PackItem, check, lib

MODE_STORE, MODE_POOL, MODE_HEAD, MODE_UNPOOL, MODE_CTX, MODE_CTX_BWD, MODE_DFV, MODE_CTX2, MODE_CTX2_BWD = range(9)

#: number of kernels of this library launched so far (bench.py reports the count per timed region)
LAUNCHES = 0
#: when a list, every tensor-core launch appends (start_event, end_event, algorithmic_flops)
PROFILE = None


def _count(n=1):
    global LAUNCHES
    LAUNCHES += n


class _Prof:
    def __init__(self, flops, label=""):
        self.flops = flops
        self.label = label

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *a):
        if PROFILE is not None:
            self.e1.record()
            PROFILE.append((self.e0, self.e1, self.flops, self.label))


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


@dataclass
class Act:
    buf: torch.Tensor  # bf16 [P, N, H, W, cs]
    C: int
    coff: int = 0

    @property
    def nsplit(self):
        return self.buf.shape[0]

    @property
    def N(self):
        return self.buf.shape[1]

    @property
    def H(self):
        return self.buf.shape[2]

    @property
    def W(self):
        return self.buf.shape[3]

    @property
    def cs(self):
        return self.buf.shape[4]

    @property
    def plane(self):
        return self.buf.stride(0)

    def view(self, C, coff):
        return Act(self.buf, C, coff)

    def image(self, n):
        """One image of the batch as its own Act (same planes, N = 1)."""
        return Act(self.buf[:, n:n + 1], self.C, self.coff)

    def to_nchw(self) -> torch.Tensor:
        out = torch.empty(self.N, self.C, self.H, self.W, device=self.buf.device, dtype=torch.float32)
        check(lib().can_nhwc_to_nchw_f32(self.buf.data_ptr(), self.plane, self.nsplit, self.cs, self.coff,
                                        out.data_ptr(), self.N, self.C, self.H, self.W, _stream()), "nhwc_to_nchw")
        return out


_RESERVED = {}   # device index -> bytes the caching allocator had reserved when a large buffer was last allocated
_COLD_READ_MIN = 512 << 20


def _read_if_fresh(t: torch.Tensor) -> None:
    """One ordinary read pass over a LARGE buffer that made the caching allocator fetch new device memory.

    On this pool's B200s the first TMA loads from feature maps of >= ~0.5 GiB whose pages no kernel had read before die,
    rarely, with an "unspecified launch failure" (`profiles/r02_k1_cold_page_fault.txt`: never on a
    second call, never after an ordinary read pass).  Memory the allocator hands out again has been read, so this runs on
    the first calls of a process at a new size only (0.2 ms per GB); not during graph capture.  CAN_B200_COLD_READ=0: off."""
    if t.numel() * t.element_size() < _COLD_READ_MIN or not t.is_cuda or os.environ.get("CAN_B200_COLD_READ", "1") != "1":
        return
    if torch.cuda.is_current_stream_capturing():
        return
    idx = t.device.index if t.device.index is not None else torch.cuda.current_device()
    reserved = torch.cuda.memory_reserved(idx)
    if reserved > _RESERVED.get(idx, 0):
        _RESERVED[idx] = reserved
        t.view(-1).view(torch.int32).sum()   # (an ATen reduction: plain loads of every byte; the value is dropped)


def new_act(nsplit, N, H, W, C, device, cs=None, zero=False) -> Act:
    cs = cs or C
    f = torch.zeros if zero else torch.empty
    buf = f(nsplit, N, H, W, cs, device=device, dtype=torch.bfloat16)
    _read_if_fresh(buf)
    return Act(buf, C, 0)


def act_from_nchw(x: torch.Tensor, nsplit: int) -> Act:
    N, Cc, H, W = x.shape
    a = new_act(nsplit, N, H, W, Cc, x.device)
    x = x.contiguous().float()
    check(lib().can_nchw_f32_to_nhwc(x.data_ptr(), a.buf.data_ptr(), a.plane, nsplit, a.cs, 0, N, Cc, H, W, _stream()),
          "nchw_to_nhwc")
    return a


# ------------------------------------------------------------------------------ weights
def pack_conv_weight(w: torch.Tensor, nsplit: int, dgrad: bool = False) -> torch.Tensor:
    """OIHW fp32 -> [plane][Cout][taps][Cin] bf16 (dgrad: [plane][Cin][taps flipped][Cout])."""
    co, ci, kh, kw = w.shape
    taps = kh * kw
    w = w.detach().contiguous().float()
    rows, k = (ci, taps * co) if dgrad else (co, taps * ci)
    dst = torch.empty(nsplit, rows, k, device=w.device, dtype=torch.bfloat16)
    check(lib().can_pack_conv_weight(w.data_ptr(), dst.data_ptr(), co, ci, taps, nsplit, int(dgrad), _stream()),
          "pack_conv_weight")
    _count()
    return dst


def pack_first_weight(w: torch.Tensor) -> torch.Tensor:
    """frontend.0 weight [64,3,3,3] as a K-major [64][64] bf16 matrix (k = ci*9 + tap, zero-padded from 27)."""
    wp = torch.zeros(64, 64, 1, 1, device=w.device, dtype=torch.float32)
    wp[:, :27, 0, 0] = w.detach().float().reshape(64, 27)
    return pack_conv_weight(wp, 1)


def pack_conv_weight_both(w: torch.Tensor, nsplit: int):
    """(fwd pack, dgrad pack) of one OIHW fp32 weight in a single pass over it."""
    co, ci, kh, kw = w.shape
    taps = kh * kw
    w = w.detach().contiguous().float()
    fwd = torch.empty(nsplit, co, taps * ci, device=w.device, dtype=torch.bfloat16)
    dg = torch.empty(nsplit, ci, taps * co, device=w.device, dtype=torch.bfloat16)
    check(lib().can_pack_conv_weight_both(w.data_ptr(), fwd.data_ptr(), dg.data_ptr(), co, ci, taps, nsplit, _stream()),
          "pack_conv_weight_both")
    _count()
    return fwd, dg


def pack_conv_weights_multi(ws, nsplit: int):
    """[(fwd pack, dgrad pack)] of up to 24 OIHW fp32 conv weights (3x3 or 1x1) in one launch."""
    items = (PackItem * len(ws))()
    out = []
    for it, w in zip(items, ws):
        co, ci, kh, kw = w.shape
        taps = kh * kw
        w = w.detach()
        assert w.is_contiguous() and w.dtype == torch.float32
        fwd = torch.empty(nsplit, co, taps * ci, device=w.device, dtype=torch.bfloat16)
        dg = torch.empty(nsplit, ci, taps * co, device=w.device, dtype=torch.bfloat16)
        it.w, it.dst_fwd, it.dst_dgrad, it.cout, it.cin, it.taps = w.data_ptr(), fwd.data_ptr(), dg.data_ptr(), co, ci, taps
        out.append((fwd, dg))
    check(lib().can_pack_conv_weights_multi(items, len(ws), nsplit, _stream()), "pack_conv_weights_multi")
    _count()
    return out


def pack_ctx_w2(ws, nsplit: int) -> torch.Tensor:
    ws = [w.detach().contiguous().float() for w in ws]
    dst = torch.empty(nsplit, 2048, 512, device=ws[0].device, dtype=torch.bfloat16)
    check(lib().can_pack_ctx_w2(*[w.data_ptr() for w in ws], dst.data_ptr(), nsplit, _stream()), "pack_ctx_w2")
    _count()
    return dst


def stack_1x1(ws) -> torch.Tensor:
    """[4] x [512(co),512(ci),1,1] -> [4][co][ci] fp32 (weights as stored)."""
    return torch.stack([w.detach().float().reshape(512, 512) for w in ws]).contiguous()


def transpose_stack(ws) -> torch.Tensor:
    """[n] x [512(co),512(ci),1,1] -> [n][ci][co] fp32, one launch for all n <= 16 matrices."""
    import ctypes as C
    out = torch.empty(len(ws), 512, 512, device=ws[0].device, dtype=torch.float32)
    keep = [w.detach().contiguous().float() for w in ws]
    ptrs = (C.c_void_p * len(keep))(*[w.data_ptr() for w in keep])
    check(lib().can_transpose_f32_multi(ptrs, out.data_ptr(), len(keep), 512, 512, _stream()), "transpose_f32_multi")
    _count()
    return out


# ------------------------------------------------------------------------------ forward
def conv_first(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, nsplit: int) -> Act:
    N, _, H, W = x.shape
    out = new_act(nsplit, N, H, W, 64, x.device)
    check(lib().can_conv_first_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), out.buf.data_ptr(), out.plane, nsplit,
                                   N, H, W, _stream()), "conv_first_fwd")
    _count()
    return out


def conv_tc(x: Act, wp: torch.Tensor, Cout: int, *, taps=9, dil=1, mode=MODE_STORE, relu=True, bias=None, out: Act = None,
            out_f32=None, mask: Act = None, arg=None, head_w=None, head_b=None, head_out=None, tabA=None, tabB=None,
            OH=0, OW=0, bn=0, stages=0, add: Act = None, ctx_wint=None, ctx_ttab=None, db=None):
    a = ConvArgs()
    a.N, a.H, a.W = x.N, x.H, x.W
    a.Cin, a.Cout = x.C, Cout
    a.taps, a.dil, a.nsplit, a.mode, a.relu, a.bn, a.stages = taps, dil, x.nsplit, mode, int(relu), bn, stages
    a.x, a.x_plane, a.x_cs, a.x_coff = x.buf.data_ptr(), x.plane, x.cs, x.coff
    a.w = wp.data_ptr()
    a.bias = _ptr(bias)
    if out is not None:
        a.out, a.out_plane, a.out_cs, a.out_coff = out.buf.data_ptr(), out.plane, out.cs, out.coff
    a.out_f32 = _ptr(out_f32)
    if mask is not None:
        a.mask, a.mask_cs, a.mask_coff, a.mask_plane = mask.buf.data_ptr(), mask.cs, mask.coff, mask.plane
    if add is not None:
        a.add, a.add_plane, a.add_cs, a.add_coff = add.buf.data_ptr(), add.plane, add.cs, add.coff
    if arg is not None:
        a.arg, a.arg_cs = arg.data_ptr(), arg.shape[-1]
    a.OH, a.OW = OH, OW
    a.head_w, a.head_b, a.head_out = _ptr(head_w), _ptr(head_b), _ptr(head_out)
    a.tabA, a.tabB = _ptr(tabA), _ptr(tabB)
    a.ctx_wint, a.ctx_ttab = _ptr(ctx_wint), _ptr(ctx_ttab)
    a.db = _ptr(db)
    _count()
    label = f"conv m{mode} {x.C}->{Cout} t{taps} d{dil} {x.H}x{x.W}"
    with _Prof(2.0 * taps * x.C * Cout * x.N * x.H * x.W, label):
        check(lib().can_conv_tc(C.byref(a), _stream()), f"conv_tc[{label} N={x.N}]")


def ctx_pool(fv: Act, w1: torch.Tensor, w2: torch.Tensor):
    """K1: pooled [N,50,512], tabA = conv{S}_1(pooled), tabB = conv{S}_2(tabA); w1/w2 = stack_1x1(...)"""
    N, H, W = fv.N, fv.H, fv.W
    dev = fv.buf.device
    pooled = torch.empty(N, 50, 512, device=dev, dtype=torch.float32)
    tabA = torch.empty_like(pooled)
    tabB = torch.empty_like(pooled)
    ws = torch.empty(lib().can_ctx_pool_workspace_bytes(N, H), device=dev, dtype=torch.uint8)
    check(lib().can_ctx_pool_fwd(fv.buf.data_ptr(), fv.plane, fv.nsplit, fv.cs, fv.coff, N, H, W, w1.data_ptr(),
                                 w2.data_ptr(), pooled.data_ptr(), tabA.data_ptr(), tabB.data_ptr(), ws.data_ptr(),
                                 _stream()), "ctx_pool_fwd")
    _count(4)
    return pooled, tabA, tabB


_mse_ws = {}


def mse_count(et: torch.Tensor, gt: torch.Tensor, want_grad: bool, grad_scale: float = 1.0, loss_out: torch.Tensor = None):
    """Returns (loss[1], cnt_et[N], cnt_gt[N], grad or None); et/gt are [N,1,h,w] fp32.  ``loss_out`` (fp32 [1] on the
    same device) receives the loss instead of a fresh tensor."""
    N = et.shape[0]
    HW = et[0].numel()
    dev = et.device
    et = et.contiguous()
    gt = gt.contiguous().float()
    loss = loss_out if loss_out is not None else torch.empty(1, device=dev, dtype=torch.float32)
    ce = torch.empty(N, device=dev, dtype=torch.float32)
    cg = torch.empty(N, device=dev, dtype=torch.float32)
    grad = torch.empty_like(et) if want_grad else None
    key = (dev, N, torch.cuda.current_stream(dev).cuda_stream)   # (the kernel's counter must not be shared by two streams)
    if key not in _mse_ws:
        _mse_ws[key] = torch.zeros(lib().can_mse_count_workspace_bytes(N), device=dev, dtype=torch.uint8)
    check(lib().can_mse_count_fwd_bwd(et.data_ptr(), gt.data_ptr(), N, HW, loss.data_ptr(), ce.data_ptr(),
                                      cg.data_ptr(), _ptr(grad), grad_scale, _mse_ws[key].data_ptr(), _stream()),
          "mse_count")
    _count()
    return loss, ce, cg, grad


def zero_ranges(base: torch.Tensor, ranges):
    """base[a:a + n] = 0 for up to 32 (a, n) slices of a flat fp32 tensor, one launch."""
    offs = (C.c_int64 * len(ranges))(*[a for a, _ in ranges])
    cnts = (C.c_int32 * len(ranges))(*[n for _, n in ranges])
    check(lib().can_zero_ranges(base.data_ptr(), offs, cnts, len(ranges), _stream()), "zero_ranges")
    _count()


def sgd_momentum_step(p, g, buf, lr, momentum=0.95, grad_scale=1.0, first=False):
    check(lib().can_sgd_momentum_step(p.data_ptr(), g.data_ptr(), buf.data_ptr(), p.numel(), lr, momentum, grad_scale,
                                      int(first), _stream()), "sgd_momentum_step")
    _count()


# ------------------------------------------------------------------------------ backward
_wgrad_ws = {}
_wgrad_ws_retired = []


def _wgrad_workspace(dev, need):
    """Split-K workspace shared by conv_wgrad and ctx_ut (kernels of one stream run in order).  A regrown buffer keeps
    its predecessor alive: a captured step of a smaller shape may still replay into it."""
    ws = _wgrad_ws.get(dev)
    if ws is None or ws.numel() < need:
        if ws is not None:
            _wgrad_ws_retired.append(ws)
        ws = torch.empty(max(need, 64 << 20), device=dev, dtype=torch.uint8)
        _wgrad_ws[dev] = ws
    return ws


def conv_wgrad(dz: Act, x: Act, grad: torch.Tensor, *, taps=9, dil=1, accumulate=False, alpha=1.0):
    """grad[co][ci][kh][kw] (+)= alpha * sum_pixels dz[pix][co] * x[pix + tap][ci]   (fp32, parameter layout)."""
    N, H, W = dz.N, dz.H, dz.W
    need = lib().can_conv_wgrad_workspace_bytes(N, H, W, x.C, dz.C, taps, dz.nsplit)
    dev = dz.buf.device
    ws = _wgrad_workspace(dev, need)
    assert grad.is_contiguous() and grad.dtype == torch.float32 and grad.numel() == dz.C * x.C * taps
    _count(2)
    with _Prof(2.0 * taps * x.C * dz.C * N * H * W, f"wgrad {x.C}->{dz.C} t{taps} d{dil} {H}x{W}"):
        check(lib().can_conv_wgrad(dz.buf.data_ptr(), dz.plane, dz.cs, dz.coff, x.buf.data_ptr(), x.plane, x.cs, x.coff,
                                   dz.nsplit, N, H, W, x.C, dz.C, taps, dil, grad.data_ptr(), int(accumulate),
                                   float(alpha), ws.data_ptr(), ws.numel(), _stream()), "conv_wgrad")


def im2col_first(x: torch.Tensor, nsplit: int) -> Act:
    N, _, H, W = x.shape
    out = new_act(nsplit, N, H, W, 64, x.device)
    check(lib().can_im2col_first(x.data_ptr(), out.buf.data_ptr(), out.plane, nsplit, N, H, W, _stream()), "im2col_first")
    _count()
    return out


def conv_first_tc(x: torch.Tensor, w_packed: torch.Tensor, bias: torch.Tensor, out: Act):
    """frontend.0 + ReLU straight from the fp32 NCHW image (bf16 path; no im2col buffer)."""
    N, _, H, W = x.shape
    assert out.nsplit == 1 and out.C == 64
    _count()
    with _Prof(2.0 * 27 * 64 * N * H * W, f"conv first 3->64 {H}x{W}"):
        check(lib().can_conv_first_tc_fwd(x.data_ptr(), w_packed.data_ptr(), bias.data_ptr(), out.buf.data_ptr(), out.cs,
                                          out.coff, N, H, W, _stream()), "conv_first_tc_fwd")


_FIRST_WS = {}


def conv_first_wgrad(dz: Act, x: torch.Tensor, grad: torch.Tensor, accumulate=False, alpha=1.0):
    """grad [64,3,3,3] fp32 (+)= alpha * dW of frontend.0 from dz (NHWC bf16, one plane) and the image itself."""
    N, _, H, W = x.shape
    assert dz.nsplit == 1 and dz.C == 64 and grad.is_contiguous() and grad.numel() == 64 * 27
    dev = x.device
    ws = _FIRST_WS.get(dev)
    if ws is None:
        ws = torch.empty(lib().can_conv_first_wgrad_workspace_bytes(), device=dev, dtype=torch.uint8)
        _FIRST_WS[dev] = ws
    _count(2)
    with _Prof(2.0 * 27 * 64 * N * H * W, f"wgrad first 3->64 {H}x{W}"):
        check(lib().can_conv_first_wgrad(dz.buf.data_ptr(), dz.cs, dz.coff, x.data_ptr(), N, H, W, grad.data_ptr(),
                                         int(accumulate), float(alpha), ws.data_ptr(), ws.numel(), _stream()),
              "conv_first_wgrad")


def crowd_sample(img_u8: torch.Tensor, dmap: torch.Tensor = None, flip: bool = False, ds: int = 8):
    """model/CrowdDataset.py:38-66 for one sample on the device: img uint8 [H,W] or [H,W,3] (cuda) and the optional float32
    density map [H,W] -> (img float32 [3, H//ds*ds, W//ds*ds] normalised, dmap float32 [1, H//ds, W//ds] * ds^2 or None)."""
    if img_u8.dtype != torch.uint8 or not img_u8.is_cuda or img_u8.dim() not in (2, 3):
        raise ValueError("crowd_sample: img must be a CUDA uint8 tensor [H,W] or [H,W,3]")
    img_u8 = img_u8.contiguous()
    H, W = img_u8.shape[:2]
    Cc = 1 if img_u8.dim() == 2 else img_u8.shape[2]
    rows, cols = H // ds, W // ds
    out = torch.empty(3, rows * ds, cols * ds, device=img_u8.device, dtype=torch.float32)
    od = None
    if dmap is not None:
        if dmap.dtype != torch.float32 or tuple(dmap.shape) != (H, W) or dmap.device != img_u8.device:
            raise ValueError("crowd_sample: dmap must be float32 [H,W] on the image's device")
        dmap = dmap.contiguous()
        od = torch.empty(1, rows, cols, device=img_u8.device, dtype=torch.float32)
    check(lib().can_crowd_sample(img_u8.data_ptr(), H, W, Cc, _ptr(dmap), int(flip), ds, out.data_ptr(), _ptr(od), _stream()),
          "crowd_sample")
    _count(2 if dmap is not None else 1)
    return out, od


def density_map(points_xy: torch.Tensor, H: int, W: int) -> torch.Tensor:
    """data_preparation/k_nearest_gaussian_kernel.py:14-54 on the device: points [n,2] float64 (col,row) -> float32 [H,W]."""
    pts = points_xy.reshape(-1, 2).contiguous()
    if pts.dtype != torch.float64 or not pts.is_cuda:
        raise ValueError("density_map: points must be a CUDA float64 tensor [n, 2] of (col, row)")
    n = pts.shape[0]
    out = torch.empty(H, W, device=pts.device, dtype=torch.float32)
    ws = torch.empty(lib().can_density_map_workspace_bytes(n), device=pts.device, dtype=torch.uint8)
    check(lib().can_density_map(pts.data_ptr() if n else None, n, H, W, out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()),
          "density_map")
    _count(2 if n else 1)
    return out


def bias_grad(dz: Act, db: torch.Tensor):
    """db[c] += sum_pixels dz[pix][c]; db fp32 [C], already initialised."""
    check(lib().can_bias_grad(dz.buf.data_ptr(), dz.plane, dz.nsplit, dz.cs, dz.coff, dz.C, dz.N * dz.H * dz.W,
                              db.data_ptr(), _stream()), "bias_grad")
    _count()


def head_bwd(dout: torch.Tensor, y5: Act, head_w: torch.Tensor, d_hw: torch.Tensor, d_hb: torch.Tensor) -> Act:
    dz = new_act(y5.nsplit, y5.N, y5.H, y5.W, 64, y5.buf.device)
    assert y5.cs == 64 and y5.coff == 0
    check(lib().can_head_bwd(dout.data_ptr(), y5.buf.data_ptr(), y5.plane, y5.nsplit, head_w.data_ptr(), dz.buf.data_ptr(),
                             dz.plane, y5.N * y5.H * y5.W, d_hw.data_ptr(), d_hb.data_ptr(), _stream()), "head_bwd")
    _count()
    return dz


def ctx_wint(nsplit, N, H, W, device) -> Act:
    out = new_act(nsplit, N, H, W, 64, device)
    check(lib().can_ctx_wint(out.buf.data_ptr(), out.plane, nsplit, N, H, W, _stream()), "ctx_wint")
    _count()
    return out


def ctx_ttab(tabA, tabB):
    N = tabA.shape[0]
    T = torch.empty(2, N, 2048, 64, device=tabA.device, dtype=torch.bfloat16)
    check(lib().can_ctx_ttab(tabA.data_ptr(), tabB.data_ptr(), T.data_ptr(), N, _stream()), "ctx_ttab")
    _count()
    return T


def ctx_pint(H, W, device) -> torch.Tensor:
    """[H][W][64] bf16: -1/area of every pooled window a pixel lies in (A operand of MODE_DFV's extra k-block)."""
    out = torch.empty(H, W, 64, device=device, dtype=torch.bfloat16)
    check(lib().can_ctx_pint(out.data_ptr(), H, W, _stream()), "ctx_pint")
    _count()
    return out


def ctx_dtab(dpool: torch.Tensor) -> torch.Tensor:
    """dpool [N][50][512] fp32 -> [N][512][64] bf16 (B operand of MODE_DFV's extra k-block)."""
    N = dpool.shape[0]
    out = torch.empty(N, 512, 64, device=dpool.device, dtype=torch.bfloat16)
    check(lib().can_ctx_dtab(dpool.data_ptr(), out.data_ptr(), N, _stream()), "ctx_dtab")
    _count()
    return out


def ctx_ut(wint: Act, x: Act) -> torch.Tensor:
    """U^T x for all images in one launch (bf16 path): returns [64 pos][B][C] fp32."""
    assert wint.nsplit == 1 and x.nsplit == 1 and wint.N == 1 and wint.C == 64
    B, H, W, Cc = x.N, x.H, x.W, x.C
    dev = x.buf.device
    need = lib().can_ctx_ut_workspace_bytes(B, H, W, Cc)
    ws = _wgrad_workspace(dev, need)
    out = torch.empty(64, B, Cc, device=dev, dtype=torch.float32)
    _count(2)
    with _Prof(2.0 * 64 * Cc * B * H * W, f"wgrad ut {Cc}->64 x{B} {H}x{W}"):
        check(lib().can_ctx_ut(wint.buf.data_ptr(), wint.cs, x.buf.data_ptr(), x.cs, x.coff, B, H, W, Cc, out.data_ptr(),
                               ws.data_ptr(), ws.numel(), _stream()), "ctx_ut")
    return out


def ctx_pos_gemv(Wm, inp, in_ns, in_ps, in_js, add, N):
    out = torch.empty(N, 50, 512, device=Wm.device, dtype=torch.float32)
    check(lib().can_ctx_pos_gemv(Wm.data_ptr(), inp.data_ptr(), in_ns, in_ps, in_js, _ptr(add), out.data_ptr(), N,
                                 _stream()), "ctx_pos_gemv")
    _count()
    return out


def ctx_outer(A, a_ns, a_ps, a_js, B, N, dW, alpha=1.0, accumulate=False):
    check(lib().can_ctx_outer(A.data_ptr(), a_ns, a_ps, a_js, B.data_ptr(), N, dW.data_ptr(), float(alpha),
                              int(accumulate), _stream()), "ctx_outer")
    _count()


# ------------------------------------------------------------------------------ probes
def bilinear_taps(in_size, out_size, device):
    i0 = torch.empty(out_size, device=device, dtype=torch.int32)
    i1 = torch.empty_like(i0)
    l0 = torch.empty(out_size, device=device, dtype=torch.float32)
    l1 = torch.empty_like(l0)
    check(lib().can_bilinear_taps(in_size, out_size, i0.data_ptr(), i1.data_ptr(), l0.data_ptr(), l1.data_ptr(),
                                  _stream()), "bilinear_taps")
    return i0, i1, l0, l1


def adaptive_windows(in_size, out_size, device):
    s = torch.empty(out_size, device=device, dtype=torch.int32)
    e = torch.empty_like(s)
    check(lib().can_adaptive_windows(in_size, out_size, s.data_ptr(), e.data_ptr(), _stream()), "adaptive_windows")
    return s, e
