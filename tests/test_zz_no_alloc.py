"""Collected last on purpose (file name): the allocation-free loss/count entry point.

``can_mse_count_fwd_bwd_static`` serves hosts that may not allocate on the device while they run --
a plugin inside a server that pre-reserved the GPU and passes it only input/output buffers, or a
loop that cannot tolerate cudaMalloc's device-wide stall.  Its scratch is an array compiled into the
library, so the test calls it through the C ABI twice in a row (the array keeps the previous call's
contents) and compares every output with ``can_mse_count_fwd_bwd``, the workspace-taking entry point.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(lib, static, et, gt, ws=None):
    n, hw = et.shape[0], et[0].numel()
    dev = et.device
    loss = torch.full((1,), float("nan"), device=dev)
    ce = torch.full((n,), float("nan"), device=dev)
    cg = torch.full((n,), float("nan"), device=dev)
    grad = torch.full_like(et, float("nan"))
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        rc = lib.can_mse_count_fwd_bwd_static(et.data_ptr(), gt.data_ptr(), n, hw, loss.data_ptr(), ce.data_ptr(),
                                              cg.data_ptr(), grad.data_ptr(), 1.0, stream)
    else:
        rc = lib.can_mse_count_fwd_bwd(et.data_ptr(), gt.data_ptr(), n, hw, loss.data_ptr(), ce.data_ptr(),
                                       cg.data_ptr(), grad.data_ptr(), 1.0, ws.data_ptr(), stream)
    assert rc == 0, lib.can_last_error_string()
    torch.cuda.synchronize()
    return loss.cpu(), ce.cpu(), cg.cpu(), grad.cpu()


def test_static_scratch_entry_point_matches_workspace_entry_point():
    import can_b200
    lib = can_b200.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    assert lib.can_mse_count_static_max_images() == 1024
    g = torch.Generator().manual_seed(21)
    # two calls in a row, the second with fewer images than the first: stale partials must not leak
    for n, h, w in [(7, 96, 128), (3, 13, 17)]:
        et = torch.randn(n, 1, h, w, generator=g).to(dev)
        gt = (torch.rand(n, 1, h, w, generator=g) * 0.05).to(dev)
        ws = torch.zeros(lib.can_mse_count_workspace_bytes(n), dtype=torch.uint8, device=dev)
        ref = _call(lib, False, et, gt, ws)
        got = _call(lib, True, et, gt)
        # same kernel, same reduction order: the tolerance of tests/test_gpu_forward.py::test_mse_count_matches_oracle
        assert float(got[0]) == pytest.approx(float(ref[0]), rel=2e-6)
        assert torch.allclose(got[1], ref[1], rtol=1e-5, atol=1e-3)
        assert torch.allclose(got[2], ref[2], rtol=1e-5, atol=1e-3)
        assert torch.allclose(got[3], ref[3], rtol=1e-6, atol=1e-7)
        assert torch.isfinite(got[3]).all()


def test_static_scratch_entry_point_rejects_oversized_batches():
    import can_b200
    lib = can_b200.lib()
    dev = torch.device("cuda:0")
    limit = lib.can_mse_count_static_max_images()
    n = limit + 1
    et = torch.zeros(n, 1, 2, 2, device=dev)
    out = torch.zeros(n, device=dev)
    rc = lib.can_mse_count_fwd_bwd_static(et.data_ptr(), et.data_ptr(), n, 4, out.data_ptr(), out.data_ptr(),
                                          out.data_ptr(), None, 1.0, None)
    assert rc < 0
    msg = lib.can_last_error_string().decode()
    assert str(limit) in msg and "exceeds" in msg
