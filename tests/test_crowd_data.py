"""Data path: oracle vs the cv2/torchvision golden vectors on the CPU, CUDA kernels vs both."""
import os

import numpy as np
import pytest
import torch

from oracle import crowd_data_oracle as D

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ("rgb_odd", "rgb_flip", "gray", "mult8", "rgb_big")


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def data_golden():
    return np.load(os.path.join(HERE, "golden", "crowd_data_golden.npz"))


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_cv2_and_torchvision(data_golden, name):
    g = data_golden
    img, gt = D.preprocess(g[name + ".img"], g[name + ".gt"], int(g[name + ".flip"]))
    assert img.shape == g[name + ".out_img"].shape and gt.shape == g[name + ".out_gt"].shape
    assert np.array_equal(img, g[name + ".out_img"])                      # float64 pipeline restated exactly
    # float32 pipeline: cv2's vectorised multiply-adds differ by at most one ulp of the largest term (3.8e-6 at 40)
    assert np.allclose(gt, g[name + ".out_gt"], rtol=2e-7, atol=5e-6)


def test_coefficients_of_the_identity_and_the_edges():
    s, f = D._coeff(8, 8)
    assert list(s) == list(range(8)) and not f.any()
    s, f = D._coeff(4, 9)                    # 9 -> 4: centres at 0.625, 2.875, 5.125, 7.375
    assert list(s) == [0, 2, 5, 7] and np.allclose(f, [0.625, 0.875, 0.125, 0.375])
    s, f = D._coeff(16, 4)                   # upscale: first/last samples clamp to the border with weight 0
    assert s[0] == 0 and f[0] == 0 and s[-1] == 3 and f[-1] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_data_path_matches_reference(dev, data_golden, name):
    from can_b200 import ops
    g = data_golden
    img_u8 = torch.from_numpy(g[name + ".img"]).to(dev)
    gt = torch.from_numpy(g[name + ".gt"]).to(dev)
    flip = bool(int(g[name + ".flip"]))
    out, od = ops.crowd_sample(img_u8, gt, flip=flip)
    o_img, o_gt = D.preprocess(g[name + ".img"], g[name + ".gt"], flip)
    assert np.array_equal(out.cpu().numpy(), g[name + ".out_img"])        # bit-exact against cv2 + torchvision
    assert np.array_equal(od.cpu().numpy(), o_gt)                         # bit-exact against the oracle
    assert np.allclose(od.cpu().numpy(), g[name + ".out_gt"], rtol=2e-7, atol=5e-6)
    assert np.array_equal(out.cpu().numpy(), o_img)
    only_img, none = ops.crowd_sample(img_u8, None, flip=flip)
    assert none is None and torch.equal(only_img, out)


@pytest.mark.gpu
def test_cuda_data_path_feeds_the_model(dev):
    """768x1024 uint8 image -> normalised NCHW fp32 -> CANNet forward (shape contract of the whole chain) + argument errors."""
    from can_b200 import ops
    from can_b200._lib import CanB200Error
    g = torch.Generator().manual_seed(1)
    img = torch.randint(0, 256, (771, 1029, 3), generator=g, dtype=torch.uint8).to(dev)
    dm = torch.rand(771, 1029, generator=g).to(dev)
    x, y = ops.crowd_sample(img, dm, flip=True)
    assert x.shape == (3, 768, 1024) and y.shape == (1, 96, 128)
    ref_x, ref_y = D.preprocess(img.cpu().numpy(), dm.cpu().numpy(), True)
    assert np.array_equal(x.cpu().numpy(), ref_x) and np.array_equal(y.cpu().numpy(), ref_y)
    with pytest.raises(ValueError):
        ops.crowd_sample(img.float(), dm)
    with pytest.raises(CanB200Error):
        ops.crowd_sample(img[:4], None)      # fewer than 8 rows
