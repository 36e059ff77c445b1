for rep in 0 1; do for h in 0 1; do
  CAN_B200_HALO1=$h timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('HALO1=$h', round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['clocks']['reasons'])"
done; done
