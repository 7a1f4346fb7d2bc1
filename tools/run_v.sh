# r02v: frames in flight / wavefront halves on the final build (Python-level bench parameters only)
B="python bench.py --steps 60 --no-cpu-baseline --no-extras"
for a in "--frames-in-flight 2" "--frames-in-flight 3" "--frames-in-flight 4" "--frames-in-flight 2 --halves 2" "--frames-in-flight 3 --halves 2" "--frames-in-flight 2"; do
  $B $a 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('%-34s %.1f Mp/s %.3f ms  e2e %.1f Mp/s %.3f ms' % ('$a', d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6, d['e2e']['ms_per_step']))"
done
