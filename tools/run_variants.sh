#!/bin/bash
# Time prebuilt variants (tools/build_variants.py) on the GPU box: the C3 bench step and the traversal micro-benchmark.
# usage: tools/run_variants.sh name1 name2 ...      env TRAV=0 skips the traversal benchmark
cd "$(dirname "$0")/.."
for v in "$@"; do
  MCRT_LIB_VARIANT="$v" python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
r=d['roofline']
print('%-14s ms/step %.3f  closest0 %.3f mixed %.3f shadowN %.3f shade %.3f  Mpaths/s %.1f' % ('$v', d['ms_per_step'], r['stage_ms_per_step']['k_trace_closest'], r['stage_ms_per_step']['k_trace_mixed'], r['stage_ms_per_step']['k_trace_shadow'], r['stage_ms_per_step']['k_shade'], d['value']/1e6))"
  if [ "${TRAV:-1}" = "1" ]; then
    MCRT_LIB_VARIANT="$v" python tools/bench_traversal.py --reps 5 2>/dev/null | python -c "
import json,sys
for line in sys.stdin:
    d=json.loads(line)
    if 'set' in d: print('    %-10s closest %.0f  anyhit %.0f Mrays/s' % (d['set'], d['closest_Mrays_s'], d['anyhit_Mrays_s']))"
  fi
done
