"""torchrun --nproc-per-node 2 tools/gpu_check_dp_graph.py : data-parallel steps replayed from the segmented CUDA
graphs must produce the losses and parameters of the eager data-parallel steps (same seeds, same kernels)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import harness  # noqa: E402
from can_b200.model import CANNet  # noqa: E402
from can_b200.utils.train_eval_utils import FusedTrainer  # noqa: E402


def run(graph: bool, rank: int, dev):
    os.environ["CAN_B200_GRAPH"] = "1" if graph else "0"
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict("he"))
    m = m.to(dev)
    tr = FusedTrainer(m, lr=2e-10, momentum=0.95)
    losses = []
    for step in range(5):
        x = harness.make_input(2, 96, 128, seed=10 * step + rank).to(dev)
        gt = harness.make_gt(2, 96, 128, seed=100 + 10 * step + rank).to(dev)
        losses.append(float(tr.step(x, gt)))
    return losses, tr.arena.clone()


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    le, ae = run(False, rank, dev)
    lg, ag = run(True, rank, dev)
    same_loss = all(abs(a - b) <= 1e-6 * abs(a) for a, b in zip(le, lg))
    dmax = (ae - ag).abs().max().item()
    other = ag.clone()
    dist.broadcast(other, src=0)
    print(f"rank {rank}: eager {le}\nrank {rank}: graph {lg}\nrank {rank}: losses equal {same_loss}, max |param diff| {dmax:.3e}, "
          f"ranks in sync {bool((other == ag).all())}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
