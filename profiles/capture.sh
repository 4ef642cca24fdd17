#!/usr/bin/env bash
# Profiling recipe of this repo (run on the B200 box through gpurun; see B200_PROFILING.md):
#   gpurun --timeout 1500 -- 'bash profiles/capture.sh r10'
# 1. the plain bench (its JSON line is the only number that counts),
# 2. the ncu launch list of the same transform legs (cold-cache, serialised: compare shares),
# 3. one `ncu --set full` capture per hot kernel.
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --legs n,tm,r,xf,next"

python bench.py --steps 10 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 600 $OUT/bench_$TAG.err

$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"

cap() {  # name regex skip
  ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 \
      -f -o $OUT/prof_${TAG}_$1 $CMD > $OUT/ncu_$1_$TAG.log 2>&1
  echo "capture $1 rc=$?"
}
cap melspec_ts melspec_ts_kernel 3
cap stft_hms50s stft_image_kernel 3
cap stft_diffeeg stft_image_kernel 8
cap eeg_transform eeg_transform_lane_kernel 3
cap gather gather_rows_kernel 3
cap spec_condition spec_condition_kernel 20        # image path, 1 024 records (launches 18-23 of the next-rows leg)
cap spec_condition_knots spec_condition_kernel 26  # fused with compute_stft, 1 024 records (launches 24-29)
cap cls_eeg_first classifier_eeg_first_kernel 2
cap cls_eeg_second classifier_eeg_second_kernel 2
cap cls_spec_column cls_spec_column_kernel 2
cap cls_spec_row cls_spec_row_kernel 2
ls -la $OUT
