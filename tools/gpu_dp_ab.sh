#!/bin/bash
# 2-GPU box: bench.py at N = 2 with 0 / 4 SMs left to NCCL, interleaved and repeated (the boards are power-capped:
# single runs differ by +-2 % with the thermal state).
OUT=${1:-gpurun_out}
export CAN_B200_BENCH_FP32=0 CAN_B200_BENCH_PARITY=0
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
python bench.py --no-cpu-baseline --steps 40 > $OUT/r02_dp2_n1_a.json 2>/dev/null
for i in 1 2 3; do for R in 0 4; do
  CAN_B200_DP_RESERVE_SMS=$R $TR bench.py --gpus 2 --steps 40 > $OUT/r02_dp2_res${R}_$i.json 2>/dev/null
done; done
python bench.py --no-cpu-baseline --steps 40 > $OUT/r02_dp2_n1_b.json 2>/dev/null
