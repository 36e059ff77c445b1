"""Checkpoint compatibility for the drop-in ``CANNet``.

The reference saves ``model.state_dict()`` of the DistributedDataParallel wrapper's *inner* module in
``train.py:158-162`` only on the non-DDP code path; checkpoints written from a wrapped model carry a
``module.`` prefix on all 42 keys, and ``train.py:100`` / ``test.py:19`` then load them with
``strict=False`` (silently matching nothing) or ``strict=True`` (failing).  These helpers accept both key
styles, verify the 42-key layout, and reproduce the positional VGG-16 graft of ``model/CANNet.py:26-35``
from a *local* torchvision-style weights file (there is no network on the training boxes).

Nothing here touches the GPU path: tensors are copied into the module's existing parameters, so a
``FusedTrainer`` arena (parameters living in one flat buffer) stays intact.
"""
from __future__ import annotations

import collections
from typing import Mapping

import torch

PREFIX = "module."

#: the 42 keys of the reference's state_dict, in order (model/CANNet.py:10-25)
FRONTEND_IDX = (0, 2, 5, 7, 10, 12, 14, 17, 19, 21)
BACKEND_IDX = (0, 2, 4, 6, 8, 10)
CTX_NAMES = ("conv1_1", "conv1_2", "conv2_1", "conv2_2", "conv3_1", "conv3_2", "conv6_1", "conv6_2")


def expected_keys():
    keys = [f"frontend.{i}.{t}" for i in FRONTEND_IDX for t in ("weight", "bias")]
    keys += [f"backend.{i}.{t}" for i in BACKEND_IDX for t in ("weight", "bias")]
    keys += ["output_layer.weight", "output_layer.bias"]
    keys += [f"{n}.weight" for n in CTX_NAMES]
    return keys


def strip_prefix(state: Mapping[str, torch.Tensor]) -> "collections.OrderedDict[str, torch.Tensor]":
    """Keys without the DistributedDataParallel ``module.`` prefix (idempotent; mixed styles are an error)."""
    keys = list(state.keys())
    n_pref = sum(k.startswith(PREFIX) for k in keys)
    if n_pref not in (0, len(keys)):
        raise ValueError(f"checkpoint mixes {n_pref} '{PREFIX}'-prefixed keys with {len(keys) - n_pref} plain ones")
    return collections.OrderedDict((k[len(PREFIX):] if n_pref else k, v) for k, v in state.items())


def add_prefix(state: Mapping[str, torch.Tensor]) -> "collections.OrderedDict[str, torch.Tensor]":
    return collections.OrderedDict((PREFIX + k, v) for k, v in strip_prefix(state).items())


def unwrap(obj):
    """Accepts a bare state_dict or the common wrappers {'state_dict': ...} / {'model': ...}."""
    if isinstance(obj, Mapping):
        for name in ("state_dict", "model", "model_state_dict"):
            if name in obj and isinstance(obj[name], Mapping):
                return obj[name]
    return obj


def load_checkpoint(model: torch.nn.Module, source, strict: bool = True, map_location="cpu"):
    """Loads a CAN checkpoint (path or mapping), with or without the ``module.`` prefix, into ``model`` in place.

    strict=True: exactly the 42 reference keys with the reference shapes, else ``KeyError`` / ``ValueError``
    naming the offending keys (the reference's ``strict=False`` at train.py:100 would load nothing and say nothing).
    strict=False: loads the intersection and returns (missing, unexpected) like ``nn.Module.load_state_dict``."""
    state = torch.load(source, map_location=map_location) if isinstance(source, (str, bytes)) or hasattr(source, "read") else source
    state = strip_prefix(unwrap(state))
    own = model.state_dict()
    missing = [k for k in own if k not in state]
    unexpected = [k for k in state if k not in own]
    if strict and (missing or unexpected):
        raise KeyError(f"checkpoint does not match the 42-key CANNet layout: missing {missing[:4]}{'...' if len(missing) > 4 else ''}, "
                       f"unexpected {unexpected[:4]}{'...' if len(unexpected) > 4 else ''}")
    bad = [k for k in own if k in state and tuple(state[k].shape) != tuple(own[k].shape)]
    if bad:
        raise ValueError("shape mismatch for " + ", ".join(f"{k}: {tuple(state[k].shape)} vs {tuple(own[k].shape)}" for k in bad[:4]))
    with torch.no_grad():
        for k, p in own.items():
            if k in state:
                p.copy_(state[k].to(device=p.device, dtype=p.dtype))   # in place: arenas / DDP buckets keep their storage
    cache = getattr(model, "_wcache", None)
    if cache is not None:
        cache.invalidate()   # packed kernel-layout copies are stale now
    return missing, unexpected


def save_checkpoint(model: torch.nn.Module, path, ddp_prefix: bool = False):
    """Writes the 42-key fp32 state_dict (CPU tensors), optionally with the ``module.`` prefix some tools expect."""
    state = collections.OrderedDict((k, v.detach().float().cpu().clone()) for k, v in strip_prefix(model.state_dict()).items())
    if list(state.keys()) != expected_keys():
        raise KeyError("model does not expose the reference's 42-key layout")
    torch.save(add_prefix(state) if ddp_prefix else state, path)


def graft_vgg16(model: torch.nn.Module, vgg_state, map_location="cpu"):
    """model/CANNet.py:30-35 offline: the first 20 tensors of a torchvision VGG-16 state_dict (``features.*`` conv
    weights and biases, in order) are copied positionally into ``frontend``.  ``vgg_state``: path or mapping."""
    state = torch.load(vgg_state, map_location=map_location) if isinstance(vgg_state, (str, bytes)) or hasattr(vgg_state, "read") else vgg_state
    src = list(unwrap(state).items())
    dst = list(model.frontend.state_dict().items())
    if len(src) < len(dst):
        raise ValueError(f"VGG-16 weights hold {len(src)} tensors, the frontend needs {len(dst)}")
    with torch.no_grad():
        for (dk, dp), (sk, sv) in zip(dst, src):
            if tuple(dp.shape) != tuple(sv.shape):
                raise ValueError(f"positional graft: frontend.{dk} {tuple(dp.shape)} vs {sk} {tuple(sv.shape)}")
            dp.copy_(sv.to(device=dp.device, dtype=dp.dtype))
    cache = getattr(model, "_wcache", None)
    if cache is not None:
        cache.invalidate()
