#!/bin/bash
# Round-end evidence on one GPU box: the GPU test suite, smoke(), the bench lines (train / inference), the launch list of
# one inference step and ncu --set full of the kernels that changed last (details + raw CSV pages; no .ncu-rep kept).
OUT=${1:-gpurun_out}
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > $OUT/final_gpu_tests.log
python __graft_entry__.py smoke 2>&1 | tail -4 > $OUT/final_smoke.log
python bench.py --steps 20 --warmup 5 2> $OUT/final_train.err | tail -1 > $OUT/final_train.json
python bench.py --workload infer --steps 20 --warmup 5 2> $OUT/final_infer.err | tail -1 > $OUT/final_infer.json
NCU="ncu --clock-control none --profile-from-start off"
$NCU --metrics gpu__time_duration.sum --csv --log-file $OUT/launches_train_b4_final.csv python tools/gpu_ncu_step.py train 4 > /dev/null 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file $OUT/launches_infer_b16_final.csv python tools/gpu_ncu_step.py infer 16 > /dev/null 2>&1
I='(\(int\))?'
cap() {  # tag, regex, launches to skip, workload, batch
  $NCU --set full --import-source on --kernel-name-base demangled -k "regex:$2" -s $3 -c 1 -f -o $OUT/prof_$1 \
      python tools/gpu_ncu_step.py $4 $5 > $OUT/ncu_$1.log 2>&1
  ncu -i $OUT/prof_$1.ncu-rep --page raw --csv > $OUT/prof_$1.raw.csv 2>/dev/null
  ncu -i $OUT/prof_$1.ncu-rep --page details > $OUT/prof_$1.details.txt 2>/dev/null
  rm -f $OUT/prof_$1.ncu-rep
}
cap fin_store512      "conv_tc_kernel<${I}256, ${I}1, ${I}0, ${I}0, ${I}4, ${I}2[,>]" 3 train 4
cap fin_store512_b16  "conv_tc_kernel<${I}256, ${I}1, ${I}0, ${I}0, ${I}4, ${I}2[,>]" 3 infer 16
cap fin_wgrad_hx      "wgrad_tc_kernel<${I}128, ${I}1, ${I}3, ${I}32, ${I}1[,>]" 1 train 4
ls -la $OUT/prof_fin_* $OUT/final_*
