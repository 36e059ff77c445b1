"""CPU restatement of the reference's ground-truth density generator (TEST INFRASTRUCTURE ONLY).

Follows ``data_preparation/k_nearest_gaussian_kernel.py:14-54``:
  :37-39  distances to the 3 nearest other heads (scipy KDTree.query(points, k=4): self + 3)
  :43-47  a unit impulse at (int(row), int(col)); heads past the far edges are skipped, negative indices wrap (numpy)
  :48-51  sigma = 0.1 * (d1 + d2 + d3)   (one head only: sigma = mean(image shape) / 4)
  :52     density += scipy.ndimage.gaussian_filter(impulse, sigma, mode='constant')   (float32 image, heads in order)

scipy's gaussian_filter (third-party, scipy 1.18.1 in the build container; not under /root/reference) is restated from
its published algorithm: separable correlation with ``w[k] = exp(-0.5 k^2 / sigma^2) / sum`` for ``|k| <= int(4 sigma + 0.5)``
(``truncate = 4``), zero padding, double accumulation, the intermediate between the two passes stored in the image's
dtype (float32): contribution(y, x) = float32( float64(float32(w[y - r])) * w[x - c] ).
Two or three heads make the reference itself fail (the missing neighbours come back as inf): rejected here too.
Pinned against the real scipy calls: tests/golden/make_density_golden.py -> tests/golden/density_golden.npz.
"""
import numpy as np


def knn_sigma(points, shape):
    """points: float64 [n, 2] as (col, row).  Returns sigma per head (float64)."""
    n = len(points)
    if n == 1:
        return np.array([np.average(np.array(shape, dtype=np.float64)) / 2.0 / 2.0])
    if n < 4:
        raise ValueError("the reference is undefined for 2 or 3 heads (KDTree.query pads with inf)")
    d = np.sqrt(((points[:, None, :] - points[None, :, :]) ** 2).sum(-1))
    d.sort(axis=1)                       # column 0 is the head itself (distance 0)
    return (d[:, 1] + d[:, 2] + d[:, 3]) * 0.1


def kernel1d(sigma):
    if sigma <= 1e-15:   # scipy's gaussian_filter skips axes with sigma <= 1e-15: the impulse stays an impulse
        return np.ones(1), 0
    radius = int(4.0 * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    return phi / phi.sum(), radius


def density_map(points, shape):
    """points (col, row) float64 [n, 2]; shape (rows, cols).  float32 [rows, cols]."""
    H, W = shape
    density = np.zeros((H, W), dtype=np.float32)
    points = np.asarray(points, dtype=np.float64).reshape(-1, 2)
    if len(points) == 0:
        return density
    sig = knn_sigma(points, shape)
    for (c, r), s in zip(points, sig):
        ri, ci = int(r), int(c)          # truncation toward zero, as the reference's int(pt[..])
        if not (ri < H and ci < W):
            continue
        # negative indices address from the far side (numpy semantics of `pt2d[int(pt[1]), int(pt[0])] = 1.`, :44);
        # beyond that the reference raises IndexError
        if ri < -H or ci < -W:
            raise IndexError("annotation outside the image by more than its size")
        ri, ci = ri % H, ci % W
        w, R = kernel1d(s)
        y0, y1 = max(ri - R, 0), min(ri + R, H - 1)
        x0, x1 = max(ci - R, 0), min(ci + R, W - 1)
        col = w[y0 - ri + R:y1 - ri + R + 1].astype(np.float32).astype(np.float64)        # pass along rows, stored as float32
        row = w[x0 - ci + R:x1 - ci + R + 1]
        density[y0:y1 + 1, x0:x1 + 1] += (col[:, None] * row[None, :]).astype(np.float32)  # pass along columns, float32 sum
    return density
