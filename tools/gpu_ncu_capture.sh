#!/bin/bash
# Nsight Compute evidence for one training step (4 x 768x1024, bf16) and one inference step (batch 16):
#   launches_*.csv  every launch of the profiled step with its device time (shares of the step)
#   prof_<tag>.ncu-rep  --set full of one launch per kernel family named below
# Usage (GPU box, one call): bash tools/gpu_ncu_capture.sh [outdir]
OUT=${1:-gpurun_out}
set -x
python tools/gpu_ncu_step.py train 4 > $OUT/ncu_plain_train.log 2>&1 || exit 1
python tools/gpu_ncu_step.py infer 16 > $OUT/ncu_plain_infer.log 2>&1 || exit 1
NCU="ncu --clock-control none --profile-from-start off"
$NCU --metrics gpu__time_duration.sum --csv --log-file $OUT/launches_train_b4.csv python tools/gpu_ncu_step.py train 4 > /dev/null 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file $OUT/launches_infer_b16.csv python tools/gpu_ncu_step.py infer 16 > /dev/null 2>&1
rx() {  # kernel name + leading template arguments -> regex on the demangled name ("<256, 1" or "<(int)256, (int)1")
  local name=$1; shift
  local r="$name<"
  local first=1
  for a in "$@"; do
    if [ $first = 1 ]; then first=0; else r="$r, "; fi
    r="$r(\\(int\\))?$a"
  done
  echo "$r[,>]"
}
KEEP_REP="pair_store512 ctx2_bwd"   # reports kept whole (gpurun brings back at most 64 MiB); the others as CSV pages
cap() {  # tag, regex, launches to skip inside the profiled step, workload, batch
  $NCU --set full --import-source on --kernel-name-base demangled -k "regex:$2" -s $3 -c 1 -f -o $OUT/prof_$1 \
      python tools/gpu_ncu_step.py $4 $5 > $OUT/ncu_$1.log 2>&1
  ncu -i $OUT/prof_$1.ncu-rep --page raw --csv > $OUT/prof_$1.raw.csv 2>/dev/null
  ncu -i $OUT/prof_$1.ncu-rep --page details > $OUT/prof_$1.details.txt 2>/dev/null
  case " $KEEP_REP " in *" $1 "*) ;; *) rm -f $OUT/prof_$1.ncu-rep ;; esac
}
cap pair_store512      "$(rx conv_tc_kernel 256 1 0 0 4 2)"  3 train 4    # 512->512 3x3 @96x128, halo-A CTA pairs (dominant kernel)
cap pair_store512_b16  "$(rx conv_tc_kernel 256 1 0 0 4 2)"  3 infer 16
cap unpool_halo128_64  "$(rx conv_tc_kernel 64 1 3 1 8 0)"   0 train 4    # max-unpool 128->64 @384x512 (HALO)
cap unpool_pair256_128 "$(rx conv_tc_kernel 128 1 3 0 8 2)"  0 train 4    # max-unpool 256->128 @192x256 (halo-A pairs)
cap unpool_pair512_256 "$(rx conv_tc_kernel 256 1 3 0 4 2)"  0 train 4    # max-unpool 512->256 @96x128 (halo-A pairs)
cap ctx2               "$(rx conv_tc_kernel 256 1 7)"            0 train 4
cap ctx2_bwd           "$(rx conv_tc_kernel 256 1 8)"            0 train 4
cap dfv                "$(rx conv_tc_kernel 256 1 6)"            0 train 4
cap wgrad_128_3_32     "$(rx wgrad_tc_kernel 128 1 3 32 1)"    3 train 4    # 512->512 weight gradient (halo X tile)
cap wgrad_swap64       'wgrad_tc_swap64_kernel'              0 train 4
cap k1_rows            'ctx_pool_rows_kernel_v3'             0 infer 16
ls -la $OUT/prof_* ; du -sh $OUT
