"""Torch-tensor front ends of the C-ABI entry points (device memory + streams only).

``Act`` describes an activation in the kernels' layout: bf16 ``[plane][N][H][W][cs]``
where the tensor occupies channels ``[coff, coff + C)`` of the pixel stride ``cs``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import ConvArgs, check, lib

MODE_STORE, MODE_POOL, MODE_HEAD, MODE_UNPOOL, MODE_CTX = range(5)


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

    def to_nchw(self) -> torch.Tensor:
        out = torch.empty(self.N, self.C, self.H, self.W, device=self.buf.device, dtype=torch.float32)
        check(lib().can_nhwc_to_nchw_f32(self.buf.data_ptr(), self.plane, self.nsplit, self.cs, self.coff,
                                        out.data_ptr(), self.N, self.C, self.H, self.W, _stream()), "nhwc_to_nchw")
        return out


def new_act(nsplit, N, H, W, C, device, cs=None, zero=False) -> Act:
    cs = cs or C
    f = torch.zeros if zero else torch.empty
    return Act(f(nsplit, N, H, W, cs, device=device, dtype=torch.bfloat16), C, 0)


def act_from_nchw(x: torch.Tensor, nsplit: int) -> Act:
    N, Cc, H, W = x.shape
    a = new_act(nsplit, N, H, W, Cc, x.device)
    x = x.contiguous().float()
    check(lib().can_nchw_f32_to_nhwc(x.data_ptr(), a.buf.data_ptr(), a.plane, nsplit, a.cs, 0, N, Cc, H, W, _stream()),
          "nchw_to_nhwc")
    return a


def pack_conv_weight(w: torch.Tensor, nsplit: int, dgrad: bool = False) -> torch.Tensor:
    """OIHW fp32 -> [plane][Cout][taps][Cin] bf16 (dgrad: [plane][Cin][taps flipped][Cout])."""
    co, ci, kh, kw = w.shape
    taps = kh * kw
    w = w.detach().contiguous().float()
    rows, k = (ci, taps * co) if dgrad else (co, taps * ci)
    dst = torch.empty(nsplit, rows, k, device=w.device, dtype=torch.bfloat16)
    check(lib().can_pack_conv_weight(w.data_ptr(), dst.data_ptr(), co, ci, taps, nsplit, int(dgrad), _stream()),
          "pack_conv_weight")
    return dst


def pack_ctx_w2(ws, nsplit: int) -> torch.Tensor:
    ws = [w.detach().contiguous().float() for w in ws]
    dst = torch.empty(nsplit, 2048, 512, device=ws[0].device, dtype=torch.bfloat16)
    check(lib().can_pack_ctx_w2(*[w.data_ptr() for w in ws], dst.data_ptr(), nsplit, _stream()), "pack_ctx_w2")
    return dst


def transpose_stack(ws) -> torch.Tensor:
    """[4] x [512(co),512(ci),1,1] -> [4][ci][co] fp32."""
    out = torch.empty(len(ws), 512, 512, device=ws[0].device, dtype=torch.float32)
    for j, w in enumerate(ws):
        w = w.detach().contiguous().float()
        check(lib().can_transpose_f32(w.data_ptr(), out[j].data_ptr(), 512, 512, _stream()), "transpose_f32")
    return out


def conv_first(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, nsplit: int) -> Act:
    N, _, H, W = x.shape
    out = new_act(nsplit, N, H, W, 64, x.device)
    check(lib().can_conv_first_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), out.buf.data_ptr(), out.plane, nsplit,
                                   N, H, W, _stream()), "conv_first_fwd")
    return out


def conv_tc(x: Act, wp: torch.Tensor, Cout: int, *, taps=9, dil=1, mode=MODE_STORE, relu=True, bias=None, out: Act = None,
            out_f32=None, mask: Act = None, arg=None, head_w=None, head_b=None, head_out=None, tabA=None, tabB=None,
            OH=0, OW=0, bn=0, stages=0):
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
        a.mask, a.mask_cs, a.mask_coff = mask.buf.data_ptr(), mask.cs, mask.coff
    if arg is not None:
        a.arg, a.arg_cs = arg.data_ptr(), arg.shape[-1]
    a.OH, a.OW = OH, OW
    a.head_w, a.head_b, a.head_out = _ptr(head_w), _ptr(head_b), _ptr(head_out)
    a.tabA, a.tabB = _ptr(tabA), _ptr(tabB)
    check(lib().can_conv_tc(C.byref(a), _stream()), "conv_tc")


def ctx_pool(fv: Act, w1t: torch.Tensor, w2t: torch.Tensor):
    N, H, W = fv.N, fv.H, fv.W
    dev = fv.buf.device
    pooled = torch.empty(N, 50, 512, device=dev, dtype=torch.float32)
    tabA = torch.empty_like(pooled)
    tabB = torch.empty_like(pooled)
    ws = torch.empty(lib().can_ctx_pool_workspace_bytes(N, H), device=dev, dtype=torch.uint8)
    check(lib().can_ctx_pool_fwd(fv.buf.data_ptr(), fv.plane, fv.nsplit, fv.cs, fv.coff, N, H, W, w1t.data_ptr(),
                                 w2t.data_ptr(), pooled.data_ptr(), tabA.data_ptr(), tabB.data_ptr(), ws.data_ptr(),
                                 _stream()), "ctx_pool_fwd")
    return pooled, tabA, tabB


_mse_ws = {}


def mse_count(et: torch.Tensor, gt: torch.Tensor, want_grad: bool, grad_scale: float = 1.0):
    """Returns (loss[1], cnt_et[N], cnt_gt[N], grad or None); et/gt are [N,1,h,w] fp32."""
    N = et.shape[0]
    HW = et[0].numel()
    dev = et.device
    et = et.contiguous()
    gt = gt.contiguous().float()
    loss = torch.empty(1, device=dev, dtype=torch.float32)
    ce = torch.empty(N, device=dev, dtype=torch.float32)
    cg = torch.empty(N, device=dev, dtype=torch.float32)
    grad = torch.empty_like(et) if want_grad else None
    key = (dev, N)
    if key not in _mse_ws:
        _mse_ws[key] = torch.zeros(lib().can_mse_count_workspace_bytes(N), device=dev, dtype=torch.uint8)
    check(lib().can_mse_count_fwd_bwd(et.data_ptr(), gt.data_ptr(), N, HW, loss.data_ptr(), ce.data_ptr(),
                                      cg.data_ptr(), _ptr(grad), grad_scale, _mse_ws[key].data_ptr(), _stream()),
          "mse_count")
    return loss, ce, cg, grad


def sgd_momentum_step(p, g, buf, lr, momentum=0.95, grad_scale=1.0, first=False):
    check(lib().can_sgd_momentum_step(p.data_ptr(), g.data_ptr(), buf.data_ptr(), p.numel(), lr, momentum, grad_scale,
                                      int(first), _stream()), "sgd_momentum_step")


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
