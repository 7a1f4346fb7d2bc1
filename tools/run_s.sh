# r02s: code-size variants of the split shade stage (out-of-line BSDF helpers / texture fetch)
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
$B > gpurun_out/r02s_default.json 2>/dev/null
MCRT_LIB_VARIANT=bxn $B > gpurun_out/r02s_bxn.json 2>/dev/null
MCRT_LIB_VARIANT=texn $B > gpurun_out/r02s_texn.json 2>/dev/null
for f in gpurun_out/r02s_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
