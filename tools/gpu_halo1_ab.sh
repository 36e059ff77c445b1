for h in 0 1; do
  for mode in store pool dgrad; do CAN_B200_HALO1=$h timeout 30 python tools/gpu_bench_conv.py 4 768 1024 64 64 1 $mode | sed "s/^/HALO1=$h /"; done
  CAN_B200_HALO1=$h timeout 30 python tools/gpu_bench_conv.py 4 384 512 64 128 1 store | sed "s/^/HALO1=$h /"
  CAN_B200_HALO1=$h CAN_B200_DBG=16 timeout 30 python tools/gpu_bench_conv.py 4 768 1024 64 64 1 store | sed "s/^/HALO1=$h /"
done
