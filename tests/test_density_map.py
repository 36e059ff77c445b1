"""Ground-truth density generator: oracle vs the scipy golden vectors (CPU), CUDA vs oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import density_map_oracle as M

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ("sparse", "dense", "edge", "one", "none", "wrap")


@pytest.fixture(scope="module")
def dgold():
    return np.load(os.path.join(HERE, "golden", "density_golden.npz"))


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_scipy_bit_for_bit(dgold, name):
    d = M.density_map(dgold[name + ".points"], tuple(int(v) for v in dgold[name + ".shape"]))
    assert np.array_equal(d, dgold[name + ".density"])


def test_two_or_three_heads_are_rejected_like_the_reference_fails():
    with pytest.raises(ValueError):
        M.density_map(np.array([[3.0, 4.0], [10.0, 12.0]]), (32, 32))


def test_mass_is_one_per_head_away_from_the_border():
    pts = np.array([[40.0, 40.0], [44.0, 40.0], [40.0, 45.0], [46.0, 46.0]])
    assert abs(M.density_map(pts, (96, 96)).sum() - 4.0) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_density_map_matches_oracle(dev, dgold, name):
    from can_b200 import ops
    h, w = (int(v) for v in dgold[name + ".shape"])
    pts = torch.from_numpy(dgold[name + ".points"]).to(dev)
    got = ops.density_map(pts, h, w).cpu().numpy()
    ref = dgold[name + ".density"]
    # same operations in the same order; only exp() (CUDA vs glibc, <= 1 ulp in double) and numpy's pairwise sum of the
    # kernel weights may differ -> a float32 ulp here and there
    assert np.allclose(got, ref, rtol=3e-6, atol=1e-12)
    assert abs(float(got.sum()) - float(ref.sum())) <= 1e-5 * max(1.0, float(ref.sum()))
    assert (got != ref).mean() < 0.02


@pytest.mark.gpu
def test_cuda_density_map_large_and_errors(dev):
    from can_b200 import ops
    from can_b200._lib import CanB200Error
    rng = np.random.default_rng(5)
    pts = np.stack([rng.random(600) * 1024, rng.random(600) * 768], 1)
    got = ops.density_map(torch.from_numpy(pts).to(dev), 768, 1024).cpu().numpy()
    ref = M.density_map(pts, (768, 1024))
    assert np.allclose(got, ref, rtol=3e-6, atol=1e-12)
    with pytest.raises(CanB200Error):
        ops.density_map(torch.zeros(2, 2, dtype=torch.float64, device=dev), 32, 32)
    with pytest.raises(ValueError):
        ops.density_map(torch.zeros(5, 2, device=dev), 32, 32)
