for s in 0 2 1 0 2; do
  CAN_B200_STAGE256=$s python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('STAGE256=$s', d['ms_per_step'], d['value'], d['clocks']['sm_mhz'])"
done
for s in 0 2; do
  CAN_B200_STAGE256=$s python tools/gpu_bench_conv.py 4 96 128 512 512 1 store
  CAN_B200_STAGE256=$s python tools/gpu_bench_conv.py 4 96 128 512 512 2 store
  CAN_B200_STAGE256=$s python tools/gpu_bench_conv.py 4 192 256 256 256 1 store
done
