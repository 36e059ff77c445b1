"""One eager step of the benchmarked workload for Nsight Compute (kept short: ncu replays every profiled kernel ~40x).

    python tools/gpu_ncu_step.py [train|infer] [batch]
    ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:<pattern> -c <n> \
        -o gpurun_out/prof python tools/gpu_ncu_step.py train 4

Two warm-up steps run outside the profiled range (cudaProfilerStart/Stop bracket exactly one step); CUDA graphs are off so
that every kernel is an ordinary launch with its own name.
"""
import os
import sys

os.environ["CAN_B200_GRAPH"] = "0"
os.environ["CAN_B200_GRAPH_INFER"] = "0"
import torch  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))
from oracle import harness  # noqa: E402
from can_b200.model import CANNet  # noqa: E402
from can_b200.utils.train_eval_utils import FusedTrainer  # noqa: E402


def main():
    train = (sys.argv[1] if len(sys.argv) > 1 else "train") == "train"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else (4 if train else 16)
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict("he"))
    m = m.cuda()
    x = harness.make_input(B, 768, 1024).cuda()
    gt = harness.make_gt(B, 768, 1024).cuda()
    tr = FusedTrainer(m, lr=1e-10) if train else None

    def step():
        if train:
            return tr.step(x, gt)
        with torch.no_grad():
            return m(x)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    out = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("ok", float(out.float().sum()))


if __name__ == "__main__":
    main()
