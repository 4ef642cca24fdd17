#!/usr/bin/env bash
# A/B timing of row (f)2 library variants on ONE box: tools/ab_f2.sh a.so b.so ...  (files under ab_libs/)
LIB=multimodal-brain-pattern-identification_xai_b200/lib/libeegspec.so
cp $LIB /tmp/keep.so
for rep in 1 2; do
for v in "$@"; do
  cp ab_libs/$v $LIB
  echo "== $v"; python tools/time_next.py f2 2>&1 | tail -3
done
done
cp /tmp/keep.so $LIB
