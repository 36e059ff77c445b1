#!/bin/bash
# ncu --set full of the halo-A pair kernels inside one training step (4 x 768x1024): one launch each, details + raw CSV pages.
OUT=${1:-gpurun_out}
NCU="ncu --clock-control none --profile-from-start off"
cap() {  # tag, regex, launches to skip
  $NCU --set full --import-source on --kernel-name-base demangled -k "regex:$2" -s $3 -c 1 -f -o $OUT/prof_$1 \
      python tools/gpu_ncu_step.py train 4 > $OUT/ncu_$1.log 2>&1
  ncu -i $OUT/prof_$1.ncu-rep --page raw --csv > $OUT/prof_$1.raw.csv 2>/dev/null
  ncu -i $OUT/prof_$1.ncu-rep --page details > $OUT/prof_$1.details.txt 2>/dev/null
  rm -f $OUT/prof_$1.ncu-rep
}
cap ha_pool128   'conv_tc_kernel<(\(int\))?128, (\(int\))?1, (\(int\))?1, (\(int\))?0, (\(int\))?8, (\(int\))?2[,>]' 0
cap ha_store512  'conv_tc_kernel<(\(int\))?256, (\(int\))?1, (\(int\))?0, (\(int\))?0, (\(int\))?4, (\(int\))?2[,>]' 3
cap ha_dgrad128  'conv_tc_kernel<(\(int\))?128, (\(int\))?1, (\(int\))?0, (\(int\))?0, (\(int\))?8, (\(int\))?2[,>]' 1
ls -la $OUT/prof_ha_*
