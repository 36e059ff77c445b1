"""Deterministic inputs and weights for parity tests and the bench -- TEST
INFRASTRUCTURE ONLY (see ``oracle/cannet_oracle.py`` for the import rules).

Three init regimes:
  1 "ref"     reference init, model/CANNet.py:93-101: weights N(0, 0.01^2), biases 0.
              Degenerate for the contextual module (every sigmoid is exactly 0.5).
  2 "default" PyTorch Conv2d default (kaiming-uniform a=sqrt(5), bias U(+-1/sqrt(fan_in))).
  3 "he"      He-normal weights, N(0, 0.1^2) biases; sigmoid weights span 0..1.
Weights are drawn key by key from a seeded CPU generator, so the same state_dict is
reproducible on any box with the same torch build without shipping 83 MB of floats.
"""
from __future__ import annotations

import math

import torch

from .cannet_oracle import state_dict_keys, state_dict_shapes

REGIMES = ("ref", "default", "he")


def make_state_dict(regime: str = "he", seed: int = 0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    shapes = state_dict_shapes()
    sd = {}
    for k in state_dict_keys():
        shp = shapes[k]
        if k.endswith("weight"):
            fan_in = shp[1] * shp[2] * shp[3]
            if regime == "ref":
                t = torch.randn(shp, generator=g) * 0.01
            elif regime == "default":
                bound = 1.0 / math.sqrt(fan_in)  # kaiming_uniform(a=sqrt(5)) == U(+-1/sqrt(fan_in))
                t = (torch.rand(shp, generator=g) * 2 - 1) * bound
            elif regime == "he":
                t = torch.randn(shp, generator=g) * math.sqrt(2.0 / fan_in)
            else:
                raise ValueError(regime)
        else:
            wshape = shapes[k.replace("bias", "weight")]
            fan_in = wshape[1] * wshape[2] * wshape[3]
            if regime == "ref":
                t = torch.zeros(shp)
            elif regime == "default":
                bound = 1.0 / math.sqrt(fan_in)
                t = (torch.rand(shp, generator=g) * 2 - 1) * bound
            else:
                t = torch.randn(shp, generator=g) * 0.1
        sd[k] = t.to(dtype)
    return sd


def make_input(n, h, w, seed=1234):
    """randn image (ImageNet-normalised range)."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 3, h, w, generator=g)


def make_gt(n, h, w, seed=4321):
    """Non-negative, crowd-like density target at 1/8 resolution."""
    g = torch.Generator().manual_seed(seed)
    return torch.rand(n, 1, h // 8, w // 8, generator=g) * 0.05


def rel_l2(a, b):
    a = a.double().flatten()
    b = b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def rel_count(a, b):
    sa = a.double().sum()
    sb = b.double().sum()
    return float((sa - sb).abs() / sb.abs().clamp_min(1e-300))
