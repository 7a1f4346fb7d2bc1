#!/bin/bash
# A/B of MCRT_TRAV_PREFETCH (traverse.cuh) on the GPU box: partition tail diagnostic + C3 step.
cd "$(dirname "$0")/.."
for v in 0 1 2 3; do
  MCRT_NVCC_DEFS="-DMCRT_TRAV_PREFETCH=$v" python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1 || { echo "BUILD FAILED $v"; continue; }
  echo "== MCRT_TRAV_PREFETCH=$v"
  MCRT_NVCC_DEFS="-DMCRT_TRAV_PREFETCH=$v" python tools/diag_partition_stages.py 1000 2>/dev/null | grep -E '"parts": (1|8)'
  MCRT_NVCC_DEFS="-DMCRT_TRAV_PREFETCH=$v" python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); r=d['roofline']['stage_ms_per_step']
print('C3 ms/step %.3f closest0 %.3f mixed %.3f shadowN %.3f shade %.3f' % (d['ms_per_step'], r['k_trace_closest'], r['k_trace_mixed'], r['k_trace_shadow'], r['k_shade']))"
done
python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1
