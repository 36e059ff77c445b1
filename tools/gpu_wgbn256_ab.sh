#!/bin/bash
for v in 256 128 256 128; do
  CAN_B200_WGBN256=$v python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('WGBN256=$v', round(d['ms_per_step'],3), 'ms', round(d['value'],1), 'img/s', d['clocks']['sm_mhz'])"
done
