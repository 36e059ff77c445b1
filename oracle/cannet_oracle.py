"""CPU oracle for the CANNet step -- TEST INFRASTRUCTURE ONLY.

This file is a plain numpy restatement of the arithmetic the reference performs
through PyTorch/ATen.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  The
product path (``can_b200``) never does: it fails loudly when the CUDA library
is missing.

Parity pin: the reference ships no golden vectors or tests of its own
(SURVEY.md section 8c), so this oracle is pinned against outputs of the live
reference module (``/root/reference/model/CANNet.py`` path-loaded in the build
container) committed under ``tests/golden/`` by ``tests/golden/make_golden.py``.

Reference lines restated (all in ``model/CANNet.py`` unless noted):
  make_layers            :104-121   conv3x3 (pad = dilation) + ReLU, 'M' = maxpool 2x2/2
  forward                :39-91     frontend -> 4 x (pool, 1x1, upsample, contrast, 1x1,
                                    sigmoid) -> fusion (:84) -> cat (:87) -> backend -> head
  MSELoss(sum)           utils/train_eval_utils.py:20,37
  count                  utils/train_eval_utils.py:83, test.py:32
  SGD momentum           train.py:63,125-126 (torch.optim.SGD, dampening 0, no nesterov)
Index semantics come from ATen (not vendored in the reference; torch 2.11):
  adaptive pool windows  ATen/native/AdaptivePooling.h:32-38
  bilinear align_corners ATen/native/UpSample.h:271-295,442-476
"""
from __future__ import annotations

import numpy as np

FRONTEND_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512]
BACKEND_CFG = [512, 512, 512, 256, 128, 64]
SCALES = (1, 2, 3, 6)
FUSE_EPS = 0.000000000001  # literal at CANNet.py:84

# Sequential indices that hold convolutions (CANNet.py:104-121 numbering)
FRONTEND_CONV_IDX = (0, 2, 5, 7, 10, 12, 14, 17, 19, 21)
BACKEND_CONV_IDX = (0, 2, 4, 6, 8, 10)


def state_dict_keys():
    """The 42 keys of the reference state_dict, in order."""
    keys = []
    for i in FRONTEND_CONV_IDX:
        keys += [f"frontend.{i}.weight", f"frontend.{i}.bias"]
    for i in BACKEND_CONV_IDX:
        keys += [f"backend.{i}.weight", f"backend.{i}.bias"]
    keys += ["output_layer.weight", "output_layer.bias"]
    for s in SCALES:
        keys += [f"conv{s}_1.weight", f"conv{s}_2.weight"]
    return keys


def state_dict_shapes():
    shapes = {}
    cin = 3
    for i, v in zip(FRONTEND_CONV_IDX, [c for c in FRONTEND_CFG if c != "M"]):
        shapes[f"frontend.{i}.weight"] = (v, cin, 3, 3)
        shapes[f"frontend.{i}.bias"] = (v,)
        cin = v
    cin = 1024
    for i, v in zip(BACKEND_CONV_IDX, BACKEND_CFG):
        shapes[f"backend.{i}.weight"] = (v, cin, 3, 3)
        shapes[f"backend.{i}.bias"] = (v,)
        cin = v
    shapes["output_layer.weight"] = (1, 64, 1, 1)
    shapes["output_layer.bias"] = (1,)
    for s in SCALES:
        shapes[f"conv{s}_1.weight"] = (512, 512, 1, 1)
        shapes[f"conv{s}_2.weight"] = (512, 512, 1, 1)
    return shapes


# ----------------------------------------------------------------------------
# integer index tables (must be bit-exact)
# ----------------------------------------------------------------------------
def adaptive_window(i: int, in_size: int, out_size: int):
    """[start, end) of adaptive-pool window i (AdaptivePooling.h:32-38)."""
    start = (i // out_size) * in_size + ((i % out_size) * in_size) // out_size
    end = 1 + ((i + 1) * in_size - 1) // out_size
    return start, end


def adaptive_windows(in_size: int, out_size: int):
    return [adaptive_window(i, in_size, out_size) for i in range(out_size)]


def bilinear_taps(in_size: int, out_size: int, dtype=np.float32):
    """align_corners=True source taps (UpSample.h:271-295,442-476).

    Returns (i0[int32], i1[int32], lam0, lam1) per destination index.  ATen does the
    floating arithmetic in the tensor's own type: float32 for the reference's fp32
    run (the bit-exact contract), float64 only for the fp64 "truth" runs.
    """
    f32 = np.dtype(dtype).type
    if out_size > 1:
        scale = f32(in_size - 1) / f32(out_size - 1)
    else:
        scale = f32(0)
    dst = np.arange(out_size, dtype=dtype)
    src = (scale * dst).astype(dtype)
    i0 = np.minimum(np.floor(src).astype(np.int32), in_size - 1)
    lam1 = np.clip((src - i0.astype(dtype)).astype(dtype), f32(0), f32(1))
    i1 = i0 + (i0 < in_size - 1).astype(np.int32)
    lam0 = (f32(1) - lam1).astype(dtype)
    return i0.astype(np.int32), i1.astype(np.int32), lam0, lam1


# ----------------------------------------------------------------------------
# float ops
# ----------------------------------------------------------------------------
def conv3x3(x, w, b, dilation=1):
    """x [N,C,H,W], w [K,C,3,3], b [K] -> [N,K,H,W]; padding = dilation (CANNet.py:114-115)."""
    n, c, h, wd = x.shape
    k = w.shape[0]
    d = dilation
    xp = np.zeros((n, c, h + 2 * d, wd + 2 * d), dtype=x.dtype)
    xp[:, :, d:d + h, d:d + wd] = x
    out = np.zeros((n, k, h, wd), dtype=x.dtype)
    for r in range(3):
        for s in range(3):
            patch = xp[:, :, r * d:r * d + h, s * d:s * d + wd]  # [N,C,H,W]
            out += np.einsum("kc,nchw->nkhw", w[:, :, r, s], patch, optimize=True)
    if b is not None:
        out += b.reshape(1, k, 1, 1)
    return out


def relu(x):
    return np.maximum(x, 0)


def maxpool2x2(x):
    n, c, h, w = x.shape
    h2, w2 = h // 2, w // 2
    v = x[:, :, :h2 * 2, :w2 * 2].reshape(n, c, h2, 2, w2, 2)
    return v.max(axis=(3, 5))


def adaptive_avg_pool(x, s):
    n, c, h, w = x.shape
    out = np.zeros((n, c, s, s), dtype=x.dtype)
    for i, (y0, y1) in enumerate(adaptive_windows(h, s)):
        for j, (x0, x1) in enumerate(adaptive_windows(w, s)):
            out[:, :, i, j] = x[:, :, y0:y1, x0:x1].sum(axis=(2, 3)) / ((y1 - y0) * (x1 - x0))
    return out


def pool_transpose_matrix(h: int, w: int, scales=(1, 2, 3, 6), dtype=np.float64):
    """[h][w][50]: entry (y, x, pos) = 1 / area(window pos) if pixel (y, x) lies in pooled position pos (scale-major,
    row-major inside a scale), else 0 -- the transpose of F.adaptive_avg_pool2d for the four scales of
    model/CANNet.py:42,51,60,70, i.e. d fv[y][x] = sum_pos M[y][x][pos] * d pooled[pos] (windows overlap when the size
    does not divide: AdaptivePooling.h:32-38)."""
    m = np.zeros((h, w, sum(s * s for s in scales)), dtype=dtype)
    pos = 0
    for s in scales:
        for (y0, y1) in adaptive_windows(h, s):
            for (x0, x1) in adaptive_windows(w, s):
                m[y0:y1, x0:x1, pos] = 1.0 / ((y1 - y0) * (x1 - x0))
                pos += 1
    return m


def conv1x1(x, w, b=None):
    out = np.einsum("kc,nchw->nkhw", w[:, :, 0, 0], x, optimize=True)
    if b is not None:
        out = out + b.reshape(1, -1, 1, 1)
    return out


def upsample_bilinear(x, h, w):
    n, c, ih, iw = x.shape
    dt = x.dtype
    y0, y1, ly0, ly1 = bilinear_taps(ih, h, dt)
    x0, x1, lx0, lx1 = bilinear_taps(iw, w, dt)
    ly0 = ly0.astype(dt).reshape(1, 1, h, 1)
    ly1 = ly1.astype(dt).reshape(1, 1, h, 1)
    lx0 = lx0.astype(dt).reshape(1, 1, 1, w)
    lx1 = lx1.astype(dt).reshape(1, 1, 1, w)
    top = x[:, :, y0][:, :, :, x0] * lx0 + x[:, :, y0][:, :, :, x1] * lx1
    bot = x[:, :, y1][:, :, :, x0] * lx0 + x[:, :, y1][:, :, :, x1] * lx1
    return ly0 * top + ly1 * bot


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def frontend(x, sd):
    idx = iter(FRONTEND_CONV_IDX)
    for v in FRONTEND_CFG:
        if v == "M":
            x = maxpool2x2(x)
        else:
            i = next(idx)
            x = relu(conv3x3(x, sd[f"frontend.{i}.weight"], sd[f"frontend.{i}.bias"], 1))
    return x


def contextual(fv, sd, return_parts=False):
    """CANNet.py:42-84.  Returns fi (and the per-scale parts when asked)."""
    n, c, h, w = fv.shape
    num = np.zeros_like(fv)
    den = np.zeros_like(fv)
    parts = {}
    for s in SCALES:
        ave = adaptive_avg_pool(fv, s)
        a = conv1x1(ave, sd[f"conv{s}_1.weight"])
        sj = upsample_bilinear(a, h, w)
        cj = sj - fv
        wj = sigmoid(conv1x1(cj, sd[f"conv{s}_2.weight"]))
        num += wj * sj
        den += wj
        parts[s] = dict(pool=ave, a=a, s=sj, w=wj)
    fi = num / (den + fv.dtype.type(FUSE_EPS))
    return (fi, parts) if return_parts else fi


def backend(x, sd):
    for i in BACKEND_CONV_IDX:
        x = relu(conv3x3(x, sd[f"backend.{i}.weight"], sd[f"backend.{i}.bias"], 2))
    return x


def forward(x, sd, dtype=np.float64, return_intermediates=False):
    """Whole CANNet.forward (CANNet.py:39-91) on numpy arrays."""
    sd = {k: np.asarray(v, dtype=dtype) for k, v in sd.items()}
    x = np.asarray(x, dtype=dtype)
    fv = frontend(x, sd)
    fi = contextual(fv, sd)
    cat = np.concatenate([fv, fi], axis=1)
    y = backend(cat, sd)
    out = conv1x1(y, sd["output_layer.weight"], sd["output_layer.bias"])
    if return_intermediates:
        return out, dict(fv=fv, fi=fi, backend=y)
    return out


def mse_sum(et, gt):
    """MSELoss(reduction='sum') over the whole batch tensor (train_eval_utils.py:20,37)."""
    d = np.asarray(et, dtype=np.float64) - np.asarray(gt, dtype=np.float64)
    return float((d * d).sum())


def mse_sum_grad(et, gt):
    return 2.0 * (np.asarray(et, dtype=np.float64) - np.asarray(gt, dtype=np.float64))


def counts(et):
    """Per-image density integral; the reference sums the whole batch tensor
    (train_eval_utils.py:83) which equals counts(et).sum()."""
    et = np.asarray(et, dtype=np.float64)
    return et.reshape(et.shape[0], -1).sum(axis=1)


def sgd_momentum_step(p, g, buf, lr, momentum=0.95, first_step=False):
    """torch.optim.SGD (dampening 0, no nesterov, wd 0): train.py:63,125-126."""
    if first_step or buf is None:
        buf = g.copy()
    else:
        buf = momentum * buf + g
    return p - lr * buf, buf
