"""CPU restatement of the reference's per-sample data path (TEST INFRASTRUCTURE ONLY -- never imported by the product).

Follows ``model/CrowdDataset.py:35-69`` of the reference:
  :38-39  img = plt.imread(...) / 255            (uint8 HxWx{3|-} -> float64 in [0,1])
  :41-43  grayscale -> three identical channels
  :48-50  optional horizontal flip of image and density map
  :53-58  cv2.resize(img, (cols//8*8, rows//8*8))  (INTER_LINEAR, float64: double coefficients, half-pixel centres,
          source index clamped; identical sizes are a plain copy)
  :60-62  gt = cv2.resize(gt, (cols//8, rows//8)) * 64   (float32 data: float coefficients, float arithmetic)
  :64-66  float32 tensor, transforms.functional.normalize(mean, std)

cv2's INTER_LINEAR is restated from its published algorithm (OpenCV 4.x ``resize.cpp``: ``fx = (dx + 0.5) * scale - 0.5``,
``sx = floor(fx)``, ``fx -= sx``, clamp to [0, src-1] with fx = 0 at the borders; horizontal pass then vertical pass).
Pinned against cv2 4.13 run in the build container: tests/golden/make_data_golden.py -> tests/golden/crowd_data_golden.npz.
"""
import numpy as np

MEAN = np.array([0.485, 0.456, 0.406], dtype=np.float32)
STD = np.array([0.229, 0.224, 0.225], dtype=np.float32)


def _coeff(n_dst, n_src):
    """cv2 INTER_LINEAR source index and fractional weight per destination index (double)."""
    scale = n_src / n_dst
    d = np.arange(n_dst, dtype=np.float64)
    f = (d + 0.5) * scale - 0.5
    s = np.floor(f).astype(np.int64)
    f = f - s
    lo = s < 0
    hi = s >= n_src - 1
    s = np.where(lo, 0, np.where(hi, n_src - 1, s))
    f = np.where(lo | hi, 0.0, f)
    return s, f


def resize_linear(src, dw, dh):
    """cv2.resize(src, (dw, dh)) for float64 (double pipeline) or float32 (float coefficients and arithmetic)."""
    sh, sw = src.shape[:2]
    if (sw, sh) == (dw, dh):
        return src.copy()
    xi, xa = _coeff(dw, sw)
    yi, ya = _coeff(dh, sh)
    x1 = np.minimum(xi + 1, sw - 1)
    y1 = np.minimum(yi + 1, sh - 1)
    if src.dtype == np.float32:
        xa, ya = xa.astype(np.float32), ya.astype(np.float32)
        one = np.float32(1)
    else:
        one = 1.0
    shp_x = (1, dw) + (1,) * (src.ndim - 2)
    shp_y = (dh, 1) + (1,) * (src.ndim - 2)
    rows = src[:, xi] * (one - xa).reshape(shp_x) + src[:, x1] * xa.reshape(shp_x)
    return rows[yi] * (one - ya).reshape(shp_y) + rows[y1] * ya.reshape(shp_y)


def preprocess(img_u8, gt_dmap, flip, ds=8):
    """(img uint8 HxW or HxWx3, gt float32 HxW) -> (img float32 [3, H//ds*ds, W//ds*ds] normalised, gt float32 [1, H//ds, W//ds] * ds^2)."""
    img = img_u8.astype(np.float64) / 255
    if img.ndim == 2:
        img = np.repeat(img[:, :, None], 3, axis=2)
    if flip:
        img = img[:, ::-1]
        gt_dmap = gt_dmap[:, ::-1]
    rows, cols = img.shape[0] // ds, img.shape[1] // ds
    img = resize_linear(np.ascontiguousarray(img), cols * ds, rows * ds).transpose(2, 0, 1)
    gt = resize_linear(np.ascontiguousarray(gt_dmap), cols, rows)[None] * (ds * ds)
    img = img.astype(np.float32)
    img = (img - MEAN[:, None, None]) / STD[:, None, None]
    return img.astype(np.float32), gt.astype(np.float32)
