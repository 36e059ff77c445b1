"""Multi-threaded CPU restatement of the CANNet step on torch CPU ops -- TEST
INFRASTRUCTURE ONLY (same import rules as ``oracle/cannet_oracle.py``).

The reference executes every FLOP through stock ATen CPU/GPU ops
(SURVEY.md section 2.1).  This file restates ``CANNet.forward``
(model/CANNet.py:39-91) and the step body of ``train_one_epoch``
(utils/train_eval_utils.py:28-52) as a functional program over a plain
``dict`` of tensors, so that

* autograd gives gradient truth (fp64) for the backward parity tests, and
* ``bench.py`` can time "the reference's CPU path" on the GPU box's host cores
  (``cpu_baseline.kind == "port"``): ``/root/reference`` does not exist there.

Pinned against the live reference module by ``tests/golden/make_golden.py``
(run in the build container, fixtures committed).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .cannet_oracle import (BACKEND_CONV_IDX, FRONTEND_CFG, FRONTEND_CONV_IDX,
                            FUSE_EPS, SCALES)


def forward(x: torch.Tensor, sd: dict, return_intermediates: bool = False):
    it = iter(FRONTEND_CONV_IDX)
    for v in FRONTEND_CFG:
        if v == "M":
            x = F.max_pool2d(x, 2, 2)
        else:
            i = next(it)
            x = F.relu(F.conv2d(x, sd[f"frontend.{i}.weight"], sd[f"frontend.{i}.bias"], padding=1))
    fv = x
    h, w = fv.shape[2], fv.shape[3]
    num = 0
    den = 0
    for s in SCALES:
        a = F.conv2d(F.adaptive_avg_pool2d(fv, (s, s)), sd[f"conv{s}_1.weight"])
        sj = F.interpolate(a, size=(h, w), mode="bilinear", align_corners=True)
        wj = torch.sigmoid(F.conv2d(sj - fv, sd[f"conv{s}_2.weight"]))
        num = num + wj * sj
        den = den + wj
    fi = num / (den + FUSE_EPS)
    y = torch.cat((fv, fi), 1)
    for i in BACKEND_CONV_IDX:
        y = F.relu(F.conv2d(y, sd[f"backend.{i}.weight"], sd[f"backend.{i}.bias"], padding=2, dilation=2))
    out = F.conv2d(y, sd["output_layer.weight"], sd["output_layer.bias"])
    if return_intermediates:
        return out, dict(fv=fv, fi=fi, backend=y)
    return out


def loss_and_grads(x, gt, sd, dtype=torch.float64):
    """Sum-MSE loss and d(loss)/d(param) for every key (train_eval_utils.py:37-38)."""
    p = {k: v.detach().to(dtype).clone().requires_grad_(True) for k, v in sd.items()}
    out = forward(x.to(dtype), p)
    loss = F.mse_loss(out, gt.to(dtype), reduction="sum")
    grads = torch.autograd.grad(loss, list(p.values()))
    return loss.detach(), out.detach(), dict(zip(p.keys(), grads))


def train_step(x, gt, sd, bufs, lr, momentum=0.95):
    """One optimizer step restating train_eval_utils.py:28-52 + torch.optim.SGD
    (train.py:63,125-126).  ``sd``/``bufs`` are updated in place; returns the loss."""
    p = {k: v.detach().requires_grad_(True) for k, v in sd.items()}
    out = forward(x, p)
    loss = F.mse_loss(out, gt, reduction="sum")
    grads = torch.autograd.grad(loss, list(p.values()))
    with torch.no_grad():
        for (k, v), g in zip(sd.items(), grads):
            if bufs.get(k) is None:
                bufs[k] = g.clone()
            else:
                bufs[k].mul_(momentum).add_(g)
            v.sub_(bufs[k], alpha=lr)
    return loss.detach()
