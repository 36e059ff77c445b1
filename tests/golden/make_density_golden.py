"""Golden vectors for the ground-truth density generator: the reference's own calls (scipy KDTree + gaussian_filter,
data_preparation/k_nearest_gaussian_kernel.py:14-54 restated line by line; the file itself imports h5py / matplotlib,
which the build container does not have).  Output: tests/golden/density_golden.npz"""
import os

import numpy as np
import scipy.spatial
from scipy.ndimage import gaussian_filter

HERE = os.path.dirname(os.path.abspath(__file__))


def gaussian_filter_density(img_shape, points):
    density = np.zeros(img_shape, dtype=np.float32)                      # :30
    gt_count = len(points)
    if gt_count == 0:
        return density
    tree = scipy.spatial.KDTree(points.copy(), leafsize=2048)            # :35-39
    distances, locations = tree.query(points, k=4)
    for i, pt in enumerate(points):                                      # :42-52
        pt2d = np.zeros(img_shape, dtype=np.float32)
        if int(pt[1]) < img_shape[0] and int(pt[0]) < img_shape[1]:
            pt2d[int(pt[1]), int(pt[0])] = 1.
        else:
            continue
        if gt_count > 1:
            sigma = (distances[i][1] + distances[i][2] + distances[i][3]) * 0.1
        else:
            sigma = np.average(np.array(img_shape)) / 2. / 2.
        density += gaussian_filter(pt2d, sigma, mode='constant')
    return density


def main():
    rng = np.random.default_rng(7)
    out = {}
    for name, (h, w), n in (("sparse", (96, 128), 12), ("dense", (120, 160), 150), ("edge", (64, 80), 9), ("one", (48, 64), 1),
                            ("none", (40, 40), 0)):
        pts = np.stack([rng.random(n) * (w + (4 if name == "edge" else 0)), rng.random(n) * (h + (4 if name == "edge" else 0))], 1)
        out[name + ".points"], out[name + ".shape"] = pts, np.array([h, w])
        out[name + ".density"] = gaussian_filter_density([h, w], pts)
    np.savez_compressed(os.path.join(HERE, "density_golden.npz"), **out)
    print("wrote", {k: v.shape for k, v in out.items() if k.endswith("density")})


if __name__ == "__main__":
    main()
