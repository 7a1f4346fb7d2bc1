#!/bin/bash
# Time the C3 bench step under each set of environment knobs (no rebuild).
# usage: tools/env_variant_bench.sh "MCRT_COHERENCE=0" "MCRT_COHERENCE=1" ...
cd "$(dirname "$0")/.."
for v in "$@"; do
  env $v python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
r=d['roofline']
print('%-40s ms/step %.3f  closest0 %.3f mixed %.3f shadowN %.3f shade %.3f  Mpaths/s %.1f  e2e %.1f' % ('$v', d['ms_per_step'], r['stage_ms_per_step']['k_trace_closest'], r['stage_ms_per_step']['k_trace_mixed'], r['stage_ms_per_step']['k_trace_shadow'], r['stage_ms_per_step']['k_shade'], d['value']/1e6, d['e2e']['value']/1e6))"
done
