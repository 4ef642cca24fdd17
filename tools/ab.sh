#!/usr/bin/env bash
# A/B timing of library variants on ONE box: tools/ab.sh a.so b.so ...  (files under ab_libs/)
LIB=multimodal-brain-pattern-identification_xai_b200/lib/libeegspec.so
cp $LIB /tmp/keep.so
for rep in 1 2; do
for v in "$@"; do
  cp ab_libs/$v $LIB
  python bench.py --steps 30 --warmup 5 --legs n 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$v ms_per_step %.4f' % d['ms_per_step'])"
done
done
cp /tmp/keep.so $LIB
