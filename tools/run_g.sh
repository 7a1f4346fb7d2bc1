python -m pytest tests -m gpu -q -k "not benchmarked_config and not c1_in_full and not c4_bdpt_1024" > gpurun_out/r02g_pytest.log 2>&1; tail -30 gpurun_out/r02g_pytest.log
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
$B > gpurun_out/r02g_default.json 2>gpurun_out/r02g_default.err
MCRT_COHERENCE=0 $B > gpurun_out/r02g_coh0.json 2>/dev/null
for f in gpurun_out/r02g_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
python tools/parity_report.py C1 C3 C2 > gpurun_out/r02g_parity.jsonl 2> gpurun_out/r02g_parity.err; tail -3 gpurun_out/r02g_parity.err
