python -m pytest tests -m gpu -q -x -k "not benchmarked_config and not c1_in_full and not c4_bdpt_1024" > gpurun_out/r02e_pytest.log 2>&1; tail -3 gpurun_out/r02e_pytest.log
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
$B > gpurun_out/r02e_sorted.json 2>gpurun_out/r02e_sorted.err
MCRT_COHERENCE=0 $B > gpurun_out/r02e_coh0.json 2>/dev/null
MCRT_COHERENCE=0 MCRT_TRI_RECORDS=0 $B > gpurun_out/r02e_coh0_notri.json 2>/dev/null
MCRT_TRI_RECORDS=0 $B > gpurun_out/r02e_sorted_notri.json 2>/dev/null
MCRT_LIB_VARIANT=b256 MCRT_COHERENCE=0 MCRT_SHADE_CTAS_PER_SM=2 $B > gpurun_out/r02e_b256_full.json 2>/dev/null
MCRT_LIB_VARIANT=b256 MCRT_COHERENCE=0 MCRT_SHADE_CTAS_PER_SM=1 MCRT_TRACE_CTAS_PER_SM=4 $B > gpurun_out/r02e_b256_half.json 2>/dev/null
MCRT_LIB_VARIANT=b256 MCRT_COHERENCE=0 MCRT_SHADE_CTAS_PER_SM=1 MCRT_TRACE_CTAS_PER_SM=4 $B --frames-in-flight 3 > gpurun_out/r02e_b256_half_fif3.json 2>/dev/null
MCRT_LIB_VARIANT=b256 MCRT_COHERENCE=0 MCRT_SHADE_CTAS_PER_SM=1 MCRT_TRACE_CTAS_PER_SM=6 $B > gpurun_out/r02e_b256_half_t6.json 2>/dev/null
MCRT_COHERENCE=0 $B --frames-in-flight 3 > gpurun_out/r02e_coh0_fif3.json 2>/dev/null
for f in gpurun_out/r02e_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r02e_extras.json 2> gpurun_out/r02e_extras.err; tail -3 gpurun_out/r02e_extras.err
python tools/parity_report.py C1 C3 > gpurun_out/r02e_parity.jsonl 2> gpurun_out/r02e_parity.err; tail -3 gpurun_out/r02e_parity.err
