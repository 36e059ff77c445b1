"""CPU, world_size 2, gloo: the host-side data-parallel logic (bucketed all-reduce of the flat
gradient arena, DDP-mean semantics, reduce_value) that FusedTrainer runs over NCCL on the GPU box.

Equivalence checked: W ranks x N images with lr*W, gradients averaged over
ranks, equals one rank on the W*N concatenated batch with lr, because the loss is a SUM over the
local batch.  Gradients here come from the oracle's torch-CPU port (no GPU in this container)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import cannet_torch_cpu as T
from oracle import harness


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from can_b200.utils.distributed_utils import get_rank, get_world_size, is_main_process, reduce_value
    from can_b200.utils.dp import GradBuckets
    torch.set_num_threads(2)
    assert get_world_size() == world and get_rank() == rank and is_main_process() == (rank == 0)
    # reduce_value: in-place sum, optional mean (distributed_utils.py:60-70)
    v = torch.tensor([float(rank + 1)])
    assert float(reduce_value(v.clone(), average=True)) == pytest.approx((1 + 2) / 2)
    assert float(reduce_value(v.clone(), average=False)) == pytest.approx(3.0)

    sd = harness.make_state_dict("he")
    keys = list(sd.keys())
    n_img, h, w = 1, 32, 48
    x = harness.make_input(n_img, h, w, seed=1234 + rank)
    gt = harness.make_gt(n_img, h, w, seed=4321 + rank)
    loss, _, grads = T.loss_and_grads(x, gt, sd, dtype=torch.float64)
    # flat arena in parameter order, cut into the trainer's three buckets
    sizes = [sd[k].numel() for k in keys]
    total = sum(sizes)
    arena = torch.cat([grads[k].reshape(-1) for k in keys]).clone()
    n_front, n_back = sum(sizes[:20]), sum(sizes[20:34])
    buckets = GradBuckets(arena, {"backend": (n_front, n_front + n_back), "ctx": (n_front + n_back, total),
                                  "frontend": (0, n_front)})
    for tag in ("backend", "ctx", "frontend"):   # the order backward produces them
        buckets.ready(tag)
    lsum = loss.clone().reshape(1)
    buckets.reduce_scalar(lsum)
    buckets.wait()
    assert buckets.world == world and buckets.grad_scale == 1.0 / world
    # SGD step with lr*world on the rank-mean gradient
    lr0 = 1e-3
    upd = -(lr0 * world) * buckets.grad_scale * arena
    if rank == 0:
        q.put((upd.numpy().copy(), float(lsum / world)))   # by value: the worker may exit before the parent reads
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_equal_one_rank_on_the_concatenated_batch():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    upd, mean_loss = q.get(timeout=600)
    upd = torch.from_numpy(upd)
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    # single process, both images in one batch, lr0
    sd = harness.make_state_dict("he")
    x = torch.cat([harness.make_input(1, 32, 48, seed=1234 + r) for r in range(world)])
    gt = torch.cat([harness.make_gt(1, 32, 48, seed=4321 + r) for r in range(world)])
    loss, _, grads = T.loss_and_grads(x, gt, sd, dtype=torch.float64)
    ref = -1e-3 * torch.cat([grads[k].reshape(-1) for k in sd.keys()])
    assert harness.rel_l2(upd, ref) < 1e-10
    # logged loss = mean over ranks of the local sum-losses = global sum-loss / world
    assert float(mean_loss) == pytest.approx(float(loss) / world, rel=1e-10)


def test_buckets_must_tile_the_arena():
    from can_b200.utils.dp import GradBuckets
    a = torch.zeros(10)
    GradBuckets(a, {"x": (0, 4), "y": (4, 10)})
    with pytest.raises(AssertionError):
        GradBuckets(a, {"x": (0, 4), "y": (5, 10)})
    with pytest.raises(AssertionError):
        GradBuckets(a, {"x": (0, 4), "y": (4, 9)})


# ---------------------------------------------------------------- init_distributed_mode (distributed_utils.py:7-28)
def _init_worker(rank, world, port, q):
    import argparse
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    os.environ.pop("SLURM_PROCID", None)
    from can_b200.utils import distributed_utils as du
    args = argparse.Namespace(world_size=99, dist_url="env://")   # the launcher's WORLD_SIZE wins over the CLI default
    du.init_distributed_mode(args)
    ok = (args.distributed and args.rank == rank and args.world_size == world and args.gpu == rank and
          args.dist_backend == ("nccl" if torch.cuda.is_available() else "gloo") and
          du.is_dist_avail_and_initialized() and du.get_rank() == rank and du.get_world_size() == world)
    v = torch.tensor([float(10 * (rank + 1))])
    ok = ok and float(du.reduce_value(v, average=True)) == 15.0 and float(v) == 15.0   # in place, like the reference
    du.cleanup()
    ok = ok and not du.is_dist_avail_and_initialized() and du.get_world_size() == 1 and du.get_rank() == 0
    q.put((rank, bool(ok)))


def test_init_distributed_mode_under_a_launcher_environment():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_init_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    assert got == {0: True, 1: True}


def test_init_distributed_mode_without_a_launcher(monkeypatch, capsys):
    import argparse
    from can_b200.utils import distributed_utils as du
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "SLURM_PROCID"):
        monkeypatch.delenv(k, raising=False)
    args = argparse.Namespace(world_size=4, dist_url="env://")
    du.init_distributed_mode(args)
    assert args.distributed is False and "Not using distributed mode" in capsys.readouterr().out
    assert du.get_world_size() == 1 and du.get_rank() == 0 and du.is_main_process()
    t = torch.tensor([3.0])
    assert du.reduce_value(t) is t and float(t) == 3.0


def test_slurm_placement_keeps_the_command_line_world_size(monkeypatch):
    import argparse
    from can_b200.utils import distributed_utils as du
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("SLURM_PROCID", "5")
    rank, world, local = du._placement(argparse.Namespace(world_size=8))
    assert (rank, world) == (5, 8) and local == 5 % max(torch.cuda.device_count(), 1)
