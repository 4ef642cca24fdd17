#!/usr/bin/env bash
# Quick loop for mode-N kernel work (run through gpurun): parity tests, then the headline leg only.
python -m pytest tests/test_gpu_melspec.py tests/test_gpu_sliding.py -x -q 2>&1 | tail -4
python bench.py --steps 20 --warmup 5 --legs n 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('ms_per_step', d['ms_per_step'], 'frac', d['roofline']['frac'], 'rec/s', d['value'])"
