#!/bin/bash
# Time tools/bench_configs.py configs under each set of knobs; a knob set containing MCRT_NVCC_DEFS rebuilds first.
# usage: CONFIGS="C4" tools/config_variant_bench.sh "X=1" "MCRT_NVCC_DEFS=-DBDPT_NOINLINE_EVAL=0" ...
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
  env $v python tools/bench_configs.py ${CONFIGS:-C4} --steps ${STEPS:-20} 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l)
    print('%-50s %-28s ms/step %.3f  Mpaths/s %.1f  Mrays/s %.0f' % ('$v', d['config'][:28], d['ms_per_step'], d['paths_per_s']/1e6, d['mrays_per_s']))"
done
if [ $rebuilt = 1 ]; then python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1; fi
