#!/bin/bash
# weight gradient of the 128-wide three-tap tiles: X halo tile (CAN_B200_WGHX=1, default) against three X tiles per stage
for v in 0 1 0 1; do
  CAN_B200_WGHX=$v python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('WGHX=$v', round(d['ms_per_step'],3), 'ms', round(d['value'],1), 'img/s', d['clocks']['sm_mhz'])"
done
