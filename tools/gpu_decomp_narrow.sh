for mode in store pool dgrad; do for d in 0 16 64 256; do CAN_B200_DBG=$d timeout 30 python tools/gpu_bench_conv.py 4 768 1024 64 64 1 $mode; done; done
for mode in store dgrad; do for d in 0 16 64 256; do CAN_B200_DBG=$d timeout 30 python tools/gpu_bench_conv.py 4 384 512 128 128 1 $mode; done; done
for d in 0 16; do CAN_B200_DBG=$d CAN_B200_HALO=0 timeout 30 python tools/gpu_bench_conv.py 4 768 1024 64 64 1 store; done
