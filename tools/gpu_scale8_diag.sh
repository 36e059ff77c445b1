#!/bin/bash
# 8 x B200: where does the per-GPU loss at N = 8 come from -- the exchange, or waiting for the slowest of eight
# power-capped GPUs?  Same steps with and without the all-reduces (CAN_B200_DP_NOCOMM=1), per-rank step times printed.
OUT=${1:-gpurun_out}
export CAN_B200_BENCH_FP32=0 CAN_B200_BENCH_PARITY=0
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
CAN_B200_DP_NOCOMM=1 $TR bench.py --gpus 8 > $OUT/r02_scale_n8_nocomm.json 2> $OUT/r02_scale_n8_nocomm.err
$TR bench.py --gpus 8 > $OUT/r02_scale_n8_c.json 2> $OUT/r02_scale_n8_c.err
