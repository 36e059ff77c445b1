#!/bin/bash
# MODE_DFV: pooling-transpose term as one more k-block (CAN_B200_DFVKB=1, default) against thirteen table rows per pixel in
# the epilogue
for v in 0 1 0 1; do
  CAN_B200_DFVKB=$v python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('DFVKB=$v', round(d['ms_per_step'],3), 'ms', round(d['value'],1), 'img/s')"
done
