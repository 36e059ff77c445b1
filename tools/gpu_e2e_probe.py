"""Where does the end-to-end step lose time against the resident step?  (run alone or under torchrun)

Variants of the same 20 training steps, each timed with CUDA events (max over ranks):
  resident        inputs in HBM, no read-back
  resident+read   inputs in HBM, blocking loss read every step
  h2d             pinned-host inputs through DevicePrefetcher, no read-back
  h2d+read        bench.py's e2e loop (prefetch + blocking read)
  h2d+read-late   prefetch + the loss of step i read after step i+1 has been enqueued
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "can-distributed-pytorch_b200"))

import torch
import torch.distributed as dist


def main():
    from can_b200.model import CANNet
    from can_b200.utils.train_eval_utils import DevicePrefetcher, FusedTrainer
    from oracle import harness

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, H, W = 4, 768, 1024
    model = CANNet(load_weights=True)
    model.load_state_dict(harness.make_state_dict("he"))
    model = model.to(dev)
    model.precision = "bf16"
    x_host = harness.make_input(B, H, W, seed=1234 + rank).pin_memory()
    gt_host = harness.make_gt(B, H, W, seed=4321 + rank).pin_memory()
    x_dev, gt_dev = x_host.to(dev), gt_host.to(dev)
    trainer = FusedTrainer(model, lr=1e-10 * world, momentum=0.95)
    pinned = torch.empty(2, 1, dtype=torch.float32).pin_memory()

    def host_batches(n):
        return DevicePrefetcher(((x_host, gt_host) for _ in range(n)), dev)

    def resident(n):
        for _ in range(n):
            trainer.step(x_dev, gt_dev)

    def resident_read(n):
        for _ in range(n):
            trainer.step(x_dev, gt_dev).cpu()

    def h2d(n):
        for xd, gd in host_batches(n):
            trainer.step(xd, gd)

    def h2d_read(n):
        for xd, gd in host_batches(n):
            trainer.step(xd, gd).cpu()

    def h2d_read_late(n):
        prev = None
        for i, (xd, gd) in enumerate(host_batches(n)):
            loss = trainer.step(xd, gd)
            pinned[i & 1].copy_(loss, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            if prev is not None:
                prev[0].synchronize()
                float(pinned[prev[1]])
            prev = (ev, i & 1)
        prev[0].synchronize()
        float(pinned[prev[1]])

    def timed(fn):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(steps)
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms / steps

    resident(6)
    h2d_read(2)
    for rep in range(2):
        for name, fn in (("resident", resident), ("resident+read", resident_read), ("h2d", h2d), ("h2d+read", h2d_read),
                         ("h2d+read-late", h2d_read_late)):
            ms = timed(fn)
            if rank == 0:
                print(f"rep {rep} world {world} {name:15s} {ms:7.3f} ms/step  {world * B / ms * 1e3:8.1f} img/s", flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
