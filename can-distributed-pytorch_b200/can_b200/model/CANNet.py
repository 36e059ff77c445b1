"""Drop-in ``CANNet`` for ``model/CANNet.py`` of zgzhengSEU/CAN-distributed-pytorch.

Same constructor (``CANNet(load_weights=False)``), same child modules and therefore the
same 42-key ``state_dict`` (model/CANNet.py:9-35), same ``forward(x) -> [N,1,H/8,W/8]``
(model/CANNet.py:39-91).  The children are real ``nn.Conv2d`` / ``nn.Sequential`` objects
that only HOLD the parameters (so ``_initialize_weights``, ``load_state_dict``, optimizers
and DistributedDataParallel see what they expect); their ``forward`` is never called.
All arithmetic runs in libcan_b200.so (hand-written sm_100a kernels) through one
``torch.autograd.Function``.  There is no CPU or cuDNN fallback: a CPU tensor or a missing
library raises.
"""
from __future__ import annotations

import os
import warnings

import torch
import torch.nn as nn

from .. import engine, ops
from ..utils.graph_cache import GraphCache

FRONTEND_FEAT = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512]
BACKEND_FEAT = [512, 512, 512, 256, 128, 64]


def make_layers(cfg, in_channels=3, batch_norm=False, dilation=False):
    """Parameter holders in the reference's Sequential numbering (model/CANNet.py:104-121):
    conv(k3, padding=dilation=d) [+BN] + ReLU per integer entry, MaxPool2d(2,2) per 'M'."""
    d = 2 if dilation else 1
    mods = []
    c = in_channels
    for item in cfg:
        if item == "M":
            mods.append(nn.MaxPool2d(kernel_size=2, stride=2))
            continue
        mods.append(nn.Conv2d(c, item, kernel_size=3, padding=d, dilation=d))
        if batch_norm:
            mods.append(nn.BatchNorm2d(item))
        mods.append(nn.ReLU(inplace=True))
        c = item
    return nn.Sequential(*mods)


class _CANNetFunction(torch.autograd.Function):
    """forward + backward of the whole network on the CUDA kernels."""

    @staticmethod
    def forward(ctx, model, x, *params):
        # grad mode is always off inside Function.forward; needs_input_grad says whether autograd
        # recorded this call (no_grad / eval-only callers get the activation-free path)
        need_grad = any(ctx.needs_input_grad[2:])
        out, saved = engine.forward(model, x, model.nsplit, save=need_grad)
        ctx.model = model
        ctx.saved = saved
        return out

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.saved is None:
            raise RuntimeError("CANNet backward called without saved activations")
        grads = engine.backward(ctx.model, ctx.saved, grad_out)
        ctx.saved = None
        return (None, None) + tuple(grads)


class CANNet(nn.Module):
    #: "bf16" (default) or "fp32" (split-bf16 tensor-core path, fp32 parity); overridable
    #: per instance via ``model.precision = ...`` or the CAN_B200_PRECISION environment variable.
    def __init__(self, load_weights=False):
        super().__init__()
        self.frontend_feat = list(FRONTEND_FEAT)
        self.backend_feat = list(BACKEND_FEAT)
        self.frontend = make_layers(self.frontend_feat)
        self.backend = make_layers(self.backend_feat, in_channels=1024, dilation=True)
        self.output_layer = nn.Conv2d(64, 1, kernel_size=1)
        for s in (1, 2, 3, 6):
            setattr(self, f"conv{s}_1", nn.Conv2d(512, 512, kernel_size=1, bias=False))
            setattr(self, f"conv{s}_2", nn.Conv2d(512, 512, kernel_size=1, bias=False))
        self.precision = os.environ.get("CAN_B200_PRECISION", "bf16")
        self._wcache = engine.WeightCache()
        if not load_weights:
            self._initialize_weights()
            self._graft_vgg16()

    # -- reference behaviour: model/CANNet.py:26-35 ------------------------------------
    def _graft_vgg16(self):
        """Copy the first 20 tensors of torchvision's VGG-16 into ``frontend`` by position.
        Offline (no weights reachable) the N(0, 0.01) init is kept and a warning is issued."""
        try:
            from torchvision import models
            vgg = models.vgg16(pretrained=True)
        except Exception as exc:  # no network / no torchvision
            warnings.warn(f"CANNet: VGG-16 weights unavailable ({type(exc).__name__}); keeping random init")
            return
        src = list(vgg.state_dict().values())
        own = self.frontend.state_dict()
        self.frontend.load_state_dict({k: src[i] for i, k in enumerate(own.keys())})

    def _initialize_weights(self):
        """model/CANNet.py:93-101."""
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.normal_(m.weight, std=0.01)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    @property
    def nsplit(self) -> int:
        # "fp32" = two bf16 planes per value (hi + lo, 16 mantissa bits), three tensor-core
        # products, short accumulation chains spread over several TMEM accumulators;
        # "bf16x3" = three planes / six products, kept for experiments.
        try:
            return {"bf16": 1, "fp32": 2, "bf16x3": 3}[self.precision]
        except KeyError:
            raise ValueError(f"CANNet.precision must be 'bf16', 'fp32' or 'bf16x3', got {self.precision!r}")

    def forward(self, x):
        if x.dim() != 4 or x.shape[1] != 3:
            raise ValueError(f"CANNet expects [N,3,H,W], got {tuple(x.shape)}")
        if not x.is_cuda:
            raise RuntimeError("can_b200.CANNet runs on CUDA (sm_100a) only; there is no CPU fallback")
        if x.shape[2] < 8 or x.shape[3] < 8:
            raise ValueError("CANNet needs H, W >= 8 (three 2x2 max-pools)")
        p0 = self.output_layer.weight
        if p0.device != x.device:
            raise RuntimeError(f"input on {x.device} but parameters on {p0.device}")
        params = tuple(self.parameters())
        if not torch.is_grad_enabled() and x.dtype == torch.float32 and os.environ.get("CAN_B200_GRAPH_INFER", "1") != "0":
            return self._forward_no_grad(x, params)
        return _CANNetFunction.apply(self, x, *params)

    def _forward_no_grad(self, x, params):
        """Inference calls of a fixed shape replay a CUDA graph of the forward pass after two eager calls (23 launches,
        ~0.5 ms of host time per call).  Graphs are dropped as soon as any parameter changes (version counter, storage,
        or an in-place arena update announced through the weight cache); `GraphCache` bounds how many shapes keep one."""
        sig = (self._wcache.epoch, self.nsplit) + tuple((p.data_ptr(), p._version) for p in params)
        key = tuple(x.shape)
        cache = self.__dict__.get("_infer_graphs")
        if cache is None:
            cache = self.__dict__["_infer_graphs"] = GraphCache()
        ent = cache.lookup(key, sig)
        if ent is None:
            # (two eager calls of the same shape in a row first; a stream of changing image sizes never captures)
            if not cache.should_capture() or torch.cuda.is_current_stream_capturing():
                return engine.forward(self, x, self.nsplit, save=False)[0]
            try:
                xs = x.clone()
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                launches0 = ops.LAUNCHES
                with torch.cuda.graph(graph):
                    out = engine.forward(self, xs, self.nsplit, save=False)[0]
                ent = [graph, xs, out, ops.LAUNCHES - launches0]
                cache.insert(key, ent)
            except Exception as exc:  # noqa: BLE001 -- the graph is an optimisation only
                import sys
                print(f"can_b200: CUDA graph capture of the forward pass failed ({exc}); staying eager", file=sys.stderr)
                os.environ["CAN_B200_GRAPH_INFER"] = "0"
                torch.cuda.synchronize()
                return engine.forward(self, x, self.nsplit, save=False)[0]
        graph, xs, out, nk = ent
        xs.copy_(x, non_blocking=True)
        graph.replay()
        ops.LAUNCHES += nk    # kernels of one forward pass (what the replay launches)
        return out.clone()    # callers may keep results across calls; the static buffer is overwritten by the next replay
