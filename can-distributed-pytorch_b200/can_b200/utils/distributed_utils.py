"""Process-group helpers under the names the reference's training script imports from
``utils/distributed_utils.py`` (init_distributed_mode :7-28, cleanup :31-32, is_dist_avail_and_initialized
:35-41, get_world_size :44-47, get_rank :50-53, is_main_process :56-57, reduce_value :60-70): same arguments,
same fields written into ``args``, same return values.

One process per GPU, started by torchrun (or ``torch.distributed.launch --use_env``, or srun).  On a GPU box the
group is NCCL with the communicator bound to this process's device at creation; without a GPU (the CPU tests) the
same calls run over gloo.
"""
import os

import torch
import torch.distributed as dist


def _placement(args):
    """``(rank, world_size, local_device)`` handed down by the launcher, or ``None`` for a plain single process.

    torchrun exports RANK / WORLD_SIZE / LOCAL_RANK; under SLURM only the task index is known, so the world size
    stays what the command line said and tasks are dealt round-robin over the node's GPUs."""
    env = os.environ
    if "RANK" in env and "WORLD_SIZE" in env:
        return int(env["RANK"]), int(env["WORLD_SIZE"]), int(env["LOCAL_RANK"])
    if "SLURM_PROCID" in env:
        task = int(env["SLURM_PROCID"])
        return task, args.world_size, task % max(torch.cuda.device_count(), 1)
    return None


def init_distributed_mode(args):
    """Fills ``args.rank / world_size / gpu / distributed / dist_backend`` and joins the process group."""
    where = _placement(args)
    args.distributed = where is not None
    if where is None:
        print("Not using distributed mode")
        return
    args.rank, args.world_size, args.gpu = where
    on_gpu = torch.cuda.is_available()
    args.dist_backend = "nccl" if on_gpu else "gloo"
    rendezvous = getattr(args, "dist_url", "env://")
    print(f"| distributed init (rank {args.rank}): {rendezvous}", flush=True)
    bind = {}
    if on_gpu:
        torch.cuda.set_device(args.gpu)
        bind["device_id"] = torch.device("cuda", args.gpu)
    dist.init_process_group(backend=args.dist_backend, init_method=rendezvous, world_size=args.world_size,
                            rank=args.rank, **bind)
    dist.barrier()


def cleanup():
    dist.destroy_process_group()


def is_dist_avail_and_initialized():
    return dist.is_available() and dist.is_initialized()


def get_world_size():
    return dist.get_world_size() if is_dist_avail_and_initialized() else 1


def get_rank():
    return dist.get_rank() if is_dist_avail_and_initialized() else 0


def is_main_process():
    return get_rank() == 0


def reduce_value(value, average=True):
    """Sum (``average=False``) or mean of ``value`` over the ranks, written into ``value`` and returned; a single
    process gets its tensor back untouched."""
    ranks = get_world_size()
    if ranks > 1:
        with torch.no_grad():
            dist.all_reduce(value)
            if average:
                value /= ranks
    return value
