set -u
CMD="python bench.py --steps 2 --warmup 3 --legs n,r"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stft_image_kernel -s 3 -c 1 -f -o gpurun_out/prof_r3_stft_hms50s $CMD > gpurun_out/ncu_r2.log 2>&1
echo rc=$?
