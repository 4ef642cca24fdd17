set -u
OUT=gpurun_out; TAG=r12
CMD="python bench.py --steps 2 --warmup 3 --legs next"
$CMD > $OUT/plain_$TAG.log 2>&1
cap() { ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o $OUT/prof_${TAG}_$1 $CMD > $OUT/ncu_$1_$TAG.log 2>&1; echo "capture $1 rc=$?"; }
cap cls_eeg_first classifier_eeg_first_kernel 2
cap cls_eeg_second classifier_eeg_second_kernel 2
cap cls_eeg_finish classifier_eeg_finish_kernel 2
cap cls_spec_column cls_spec_column_kernel 2
