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
    # edge cases of the reference's own arithmetic: (a) slightly negative coordinates -- int() truncates toward zero, so
    # (-0.5, .) lands on index 0 and (-3.2, .) on index -3 = the third row / column from the far side (numpy indexing);
    # (b) four coincident annotations -- sigma = 0, for which scipy's gaussian_filter returns the impulse unchanged
    pts = np.array([[-0.5, 10.2], [20.7, -0.3], [-3.2, 30.0], [40.0, -2.6], [25.0, 25.0], [50.5, 12.5], [50.5, 12.5],
                    [50.5, 12.5], [50.5, 12.5], [60.0, 40.0]])
    out["wrap.points"], out["wrap.shape"] = pts, np.array([48, 64])
    out["wrap.density"] = gaussian_filter_density([48, 64], pts)
    np.savez_compressed(os.path.join(HERE, "density_golden.npz"), **out)
    print("wrote", {k: v.shape for k, v in out.items() if k.endswith("density")})


if __name__ == "__main__":
    main()
