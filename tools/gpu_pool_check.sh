for d in 0 512; do
CAN_B200_DBG=$d timeout 30 python tools/gpu_bench_conv.py 4 768 1024 64 64 1 pool
CAN_B200_DBG=$d timeout 30 python tools/gpu_bench_conv.py 4 384 512 128 128 1 pool
CAN_B200_DBG=$d timeout 30 python tools/gpu_bench_conv.py 4 192 256 256 256 1 pool
done
