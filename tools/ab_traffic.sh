#!/usr/bin/env bash
# DRAM bytes of the mode-N kernel for library variants under ab_libs/ (ncu, one launch each)
LIB=multimodal-brain-pattern-identification_xai_b200/lib/libeegspec.so
cp $LIB /tmp/keep.so
for v in "$@"; do
  cp ab_libs/$v $LIB
  echo "== $v"
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum -k regex:melspec_ts_kernel -s 3 -c 1 python bench.py --legs n --steps 2 --warmup 3 2>&1 | grep -E "dram__"
done
cp /tmp/keep.so $LIB
