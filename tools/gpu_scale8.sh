#!/bin/bash
# One 8 x B200 box: this repo's data-parallel step at N = 1 and N = 8 (bench.py, as the driver launches it) and the
# incumbent (torch eager stock ops + DistributedDataParallel, tools/gpu_incumbent.py) at N = 8.
OUT=${1:-gpurun_out}
export CAN_B200_BENCH_FP32=0 CAN_B200_BENCH_PARITY=0
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
python bench.py --no-cpu-baseline > $OUT/r02_scale_n1.json 2> $OUT/r02_scale_n1.err
$TR bench.py --gpus 8 > $OUT/r02_scale_n8.json 2> $OUT/r02_scale_n8.err
CAN_B200_DP_RESERVE_SMS=8 $TR bench.py --gpus 8 > $OUT/r02_scale_n8_res8.json 2> $OUT/r02_scale_n8_res8.err
$TR tools/gpu_incumbent.py --ddp --variants bf16_autocast_channels_last,fp32_tf32on > $OUT/r02_incumbent_ddp8.log 2>&1
$TR bench.py --gpus 8 > $OUT/r02_scale_n8_b.json 2> $OUT/r02_scale_n8_b.err
