#!/bin/bash
# Rebuild libmcrt_b200.so with each set of -D knobs ON THE GPU BOX and time the C3 bench step.
# usage: tools/variant_bench.sh "<defs1>" "<defs2>" ...   ("" = defaults)
cd "$(dirname "$0")/.."
for v in "$@"; do
  MCRT_NVCC_DEFS="$v" python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1 || { echo "BUILD FAILED: $v"; continue; }
  MCRT_NVCC_DEFS="$v" python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
r=d['roofline']
print('%-70s ms/step %.3f  closest0 %.3f mixed %.3f shadowN %.3f shade %.3f  Mpaths/s %.1f' % ('$v', d['ms_per_step'], r['stage_ms_per_step']['k_trace_closest'], r['stage_ms_per_step']['k_trace_mixed'], r['stage_ms_per_step']['k_trace_shadow'], r['stage_ms_per_step']['k_shade'], d['value']/1e6))"
done
