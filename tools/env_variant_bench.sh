#!/bin/bash
# Time the C3 bench step under each set of environment knobs; a knob set containing MCRT_NVCC_DEFS rebuilds the
# library on the GPU box first (and the last line restores the default build).
# usage: tools/env_variant_bench.sh "MCRT_COHERENCE=0" "MCRT_COHERENCE=1 MCRT_NVCC_DEFS=-DPT_SHADE_BLOCK=128" ...
cd "$(dirname "$0")/.."
rebuilt=0
for v in "$@"; do
  if [[ "$v" == *MCRT_NVCC_DEFS* ]]; then
    env $v python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1 || { echo "BUILD FAILED: $v"; continue; }
    rebuilt=1
  fi
  env $v python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
r=d['roofline']
print('%-60s ms/step %.3f  closest0 %.3f mixed %.3f shadowN %.3f shade %.3f  Mpaths/s %.1f  e2e %.1f' % ('$v', d['ms_per_step'], r['stage_ms_per_step']['k_trace_closest'], r['stage_ms_per_step']['k_trace_mixed'], r['stage_ms_per_step']['k_trace_shadow'], r['stage_ms_per_step']['k_shade'], d['value']/1e6, d['e2e']['value']/1e6))"
done
if [ $rebuilt = 1 ]; then python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1; fi
