#!/bin/bash
# epilogue of the 256-wide pair tiles: what the accumulator reads (DBG 64), the arithmetic (DBG 256 = all but the stores)
# and the stores cost; staged TMA stores (CAN_B200_STAGE256) against per-lane stores
for shape in "4 96 128 512 512 1" "4 192 256 256 256 1"; do
for m in store dgraddb; do
for d in 0 16 64 256; do CAN_B200_DBG=$d python tools/gpu_bench_conv.py $shape $m; done
for s in 1 2; do CAN_B200_STAGE256=$s python tools/gpu_bench_conv.py $shape $m | sed "s/\[/[STAGE256=$s /"; done
done
done
