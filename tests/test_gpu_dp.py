"""GPU, 2 ranks, NCCL: the data-parallel training step on hardware (SURVEY.md sections 4(v), 8e).

Replaces ``DistributedDataParallel(model, device_ids=[gpu])`` + ``args.lr *= world_size`` of the reference
(train.py:25,121-122) and the scalar loss all-reduce (utils/distributed_utils.py:60-70).  Checked on two GPUs:

* 2 ranks x 2 images with lr0 * 2 (gradients averaged over ranks)  ==  1 rank x the 4 concatenated images with lr0
  (the loss is a SUM over the local batch), on the applied parameter update and on the logged loss;
* the captured step (one CUDA graph per gradient bucket, NCCL issued between the replays) == the eager step (same
  kernels; bias gradients are accumulated with float atomics, so "equal" is to 1e-6 of the accumulated gradient);
* every rank holds bit-identical parameters after the steps.

Skipped on a box with one GPU (run with ``gpurun --gpus 2``; the log of such a run is kept under ``profiles/``).
"""
import os
import socket

import pytest
import torch

from oracle import harness

pytestmark = pytest.mark.gpu

N_PER_RANK, H, W = 2, 96, 160
LR0, STEPS = 1e-9, 4


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _batch(step, rank):
    return (harness.make_input(N_PER_RANK, H, W, seed=1000 + 10 * step + rank),
            harness.make_gt(N_PER_RANK, H, W, seed=2000 + 10 * step + rank))


def _train(dev, lr, batches, graph):
    """STEPS FusedTrainer steps on `dev`; returns (losses, flat parameter arena, momentum arena = the accumulated
    gradients 0.95^k g_k, which unlike the parameters resolve the update to fp32 precision, bucket offsets)."""
    from can_b200.model import CANNet
    from can_b200.utils.train_eval_utils import FusedTrainer
    os.environ["CAN_B200_GRAPH"] = "1" if graph else "0"
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict("he"))
    m = m.to(dev)
    m.precision = "bf16"
    tr = FusedTrainer(m, lr=lr, momentum=0.95)
    losses = [float(tr.step(x.to(dev), g.to(dev))) for x, g in batches]
    torch.cuda.synchronize(dev)
    return losses, tr.arena.clone(), tr.buf.clone(), dict(tr.buckets.ranges)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        batches = [_batch(s, rank) for s in range(STEPS)]
        # eager data-parallel steps, then the same steps with the per-bucket CUDA graphs (captured on the third step)
        l_eager, a_eager, b_eager, _ = _train(dev, LR0 * world, batches, graph=False)
        l_graph, a_graph, b_graph, _ = _train(dev, LR0 * world, batches, graph=True)
        same_as_rank0 = torch.cat([a_graph, b_graph])
        dist.broadcast(same_as_rank0, src=0)
        res = dict(rank=rank, l_eager=l_eager, l_graph=l_graph, graph_vs_eager=harness.rel_l2(b_graph, b_eager),
                   ranks_identical=bool(torch.equal(same_as_rank0, torch.cat([a_graph, b_graph]))))
        if rank == 0:
            res["buf"] = b_graph.cpu().numpy().copy()
        q.put(res)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_two_ranks_equal_one_rank_and_graph_equals_eager():
    import torch.multiprocessing as mp
    world = 2
    # one rank, the concatenated batch, lr0 (this process; no process group)
    dev = torch.device("cuda:0")
    cat = [tuple(torch.cat(t) for t in zip(*[_batch(s, r) for r in range(world)])) for s in range(STEPS)]
    l_one, _, buf_one, ranges = _train(dev, LR0, cat, graph=False)
    buf_one = buf_one.cpu().double()
    torch.cuda.empty_cache()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted((q.get(timeout=900) for _ in range(world)), key=lambda r: r["rank"])
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    for r in results:
        assert r["graph_vs_eager"] < 1e-6, "captured data-parallel step differs from the eager one"
        assert r["ranks_identical"], "ranks hold different parameters / momentum after the same steps"
        assert r["l_graph"] == pytest.approx(r["l_eager"], rel=1e-6)
        # logged loss = mean over ranks of the local sum-losses = half the sum-loss of the concatenated batch
        assert r["l_graph"] == pytest.approx([v / world for v in l_one], rel=1e-5)
    assert results[0]["l_graph"] == results[1]["l_graph"]
    # rank-mean gradients x world == gradients of the concatenated batch, bucket by bucket (fp32 summation order differs)
    buf_dp = torch.from_numpy(results[0]["buf"]).double() * world
    for tag, (a, b) in ranges.items():
        err = harness.rel_l2(buf_dp[a:b], buf_one[a:b])
        assert err < 1e-4, (tag, err)
    print("dp: losses", results[0]["l_graph"], "one rank / 2:", [v / world for v in l_one])
