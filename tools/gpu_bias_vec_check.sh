for mode in store pool dgrad; do timeout 30 python tools/gpu_bench_conv.py 4 768 1024 64 64 1 $mode; done
timeout 30 python tools/gpu_bench_conv.py 4 384 512 128 128 1 store
timeout 30 python tools/gpu_bench_conv.py 4 96 128 512 512 1 store
timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('bench', round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), round(d['roofline']['achieved'],1), d['clocks']['reasons'])"
